// epilogue.cuh -- how the result vector of a streaming pass leaves the GPU.
//
// Spark merges the partition aggregators of a pass with treeAggregate and ships the result to the driver
// (WeightedLeastSquares.fit / RDDLossFunction.calculate behind lr.fit, DataQuality4MachineLearningApp.java:126).  Here
// the last CTA of the pass -- the one that has just folded the per-CTA partials -- finishes the job inside the kernel:
//   * one GPU: it writes the vector into mapped pinned host memory and then releases a sequence number next to it; the
//     host polls that word instead of issuing a copy and a stream synchronisation;
//   * several GPUs (one process per GPU, bound with dq4ml_ctx_init_nccl): it first stores its vector into a slot of
//     EVERY rank's mailbox (plain stores to peer memory over NVLink; the mailboxes were exchanged once as CUDA IPC
//     handles), releases a per-rank flag in each mailbox, waits until its own mailbox holds all G vectors of this pass,
//     and folds them IN RANK ORDER -- so every rank computes bit-identical sums, and min / max combine as such instead of
//     riding a sum -- straight into the host buffer.  No NCCL call, no device-to-host copy and no stream synchronisation
//     are left on the per-pass critical path; messages are 48 bytes to 1 KB.
// The packed Gram of the normal-equation branch (267 KB at d = 256) keeps using ncclAllReduce.
//
// Mailbox layout (per rank, device memory): flags[2][kXrMaxRanks] (u64), then data[2][kXrMaxRanks][kXrSlotDoubles].
// Two buffers alternate with the pass number: a rank can only start pass s + 1 after it has read every slot of pass s,
// and nobody can finish pass s + 1 (and go on to overwrite buffer s % 2 with pass s + 2) before every rank has flagged
// s + 1, i.e. finished reading s.
#pragma once

#include "device_utils.cuh"
#include "kernels.hpp"

namespace dq4ml {

constexpr unsigned long long kXrErrorFlag = ~0ull;
constexpr size_t kXrFlagBytes = 2 * kXrMaxRanks * sizeof(unsigned long long);

__device__ __forceinline__ unsigned long long* xr_flag(unsigned char* box, int buf, int r) {
  return reinterpret_cast<unsigned long long*>(box) + buf * kXrMaxRanks + r;
}
__device__ __forceinline__ double* xr_slot(unsigned char* box, int buf, int r) {
  return reinterpret_cast<double*>(box + kXrFlagBytes) + (static_cast<size_t>(buf) * kXrMaxRanks + r) * kXrSlotDoubles;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double ld_relaxed_sys_f64(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Called by EVERY thread of the last CTA once thread 0 has written result[0, nacc).  kind_of(k) is the ReduceKind of
// accumulator k; `consts` are the pass's device constants (the summarizer's centres), of which ep.extra_consts are
// appended to the vector so that the host can read the shifted sums out without a second round trip.
template <class KindFn>
__device__ __forceinline__ void xr_publish(const XrEpilogue& ep, double* __restrict__ result, const int nacc,
                                           const double* __restrict__ consts, KindFn kind_of) {
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int len = nacc + ep.extra_consts;
  for (int j = tid; j < ep.extra_consts; j += nthr) result[nacc + j] = consts[j];
  if (ep.host_out == nullptr) return;  // the host fetches `result` itself (NCCL all-reduce + copy)
  __syncthreads();
  __shared__ int s_timeout;
  if (ep.world <= 1) {
    // one thread, one release: the vector is a few dozen doubles, and a single system-scope release store after the
    // writes is cheaper than a system fence in every thread
    if (tid == 0) {
      for (int k = 0; k < len; ++k) ep.host_out[k] = result[k];
      st_release_sys(ep.host_flag, ep.host_seq);
    }
    return;
  } else {
    const int buf = static_cast<int>(ep.xr_seq & 1ull);
    if (tid == 0) s_timeout = 0;
    for (int e = tid; e < ep.world * len; e += nthr) {
      const int r = e / len, k = e - r * len;
      xr_slot(ep.peer[r], buf, ep.rank)[k] = result[k];
    }
    __threadfence_system();
    __syncthreads();
    if (tid < ep.world) {
      st_release_sys(xr_flag(ep.peer[tid], buf, ep.rank), ep.xr_seq);
      const unsigned long long* mine = xr_flag(ep.peer[ep.rank], buf, tid);
      const unsigned long long t0 = global_timer_ns();
      while (ld_acquire_sys(mine) != ep.xr_seq) {
        if (global_timer_ns() - t0 > 30ull * 1000000000ull) {  // a rank died: give up instead of hanging the GPU
          s_timeout = 1;
          break;
        }
      }
    }
    __syncthreads();
    if (s_timeout != 0) {
      if (tid == 0) st_release_sys(ep.host_flag, kXrErrorFlag);
      return;
    }
    unsigned char* own = ep.peer[ep.rank];
    if (ep.gather != 0) {
      for (int e = tid; e < ep.world * len; e += nthr) {
        const int r = e / len, k = e - r * len;
        ep.host_out[e] = ld_relaxed_sys_f64(xr_slot(own, buf, r) + k);
      }
    } else {
      for (int k = tid; k < len; k += nthr) {
        double v;
        if (k < nacc) {
          const int kind = kind_of(k);
          v = ld_relaxed_sys_f64(xr_slot(own, buf, 0) + k);
          for (int r = 1; r < ep.world; ++r) v = combine(kind, v, ld_relaxed_sys_f64(xr_slot(own, buf, r) + k));
        } else {
          v = result[k];  // constants are this rank's own
        }
        ep.host_out[k] = v;
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (tid == 0) st_release_sys(ep.host_flag, ep.host_seq);
}

}  // namespace dq4ml
