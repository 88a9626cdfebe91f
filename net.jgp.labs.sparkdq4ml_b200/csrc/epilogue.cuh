// epilogue.cuh -- how the result vector of a streaming pass leaves the GPU.
//
// Spark merges the partition aggregators of a pass with treeAggregate and ships the result to the driver
// (WeightedLeastSquares.fit / RDDLossFunction.calculate behind lr.fit, DataQuality4MachineLearningApp.java:126).  Here
// the last CTA of the pass -- the one that has just folded the per-CTA partials -- finishes the job inside the kernel:
//   * one GPU: it writes the vector into mapped pinned host memory, every word tagged with the pass number; the host
//     polls those words instead of issuing a copy and a stream synchronisation;
//   * several GPUs (one process per GPU, bound with dq4ml_ctx_init_nccl): it first stores its vector into a slot of
//     EVERY rank's mailbox (plain stores to peer memory over NVLink; the mailboxes were exchanged once as CUDA IPC
//     handles), waits until its own mailbox holds all G vectors of this pass, and folds them IN RANK ORDER -- so every
//     rank computes bit-identical sums, and min / max combine as such instead of riding a sum -- straight into the host
//     buffer.  No NCCL call, no device-to-host copy, no stream synchronisation and no memory fence are left on the
//     per-pass critical path; messages are 48 bytes to 1 KB.
// The packed Gram of the normal-equation branch (267 KB at d = 256) keeps using ncclAllReduce.
//
// Wire format ("LL": low-latency, the scheme NCCL's LL protocol uses): a double travels as two 8-byte words
// {low 32 bits, tag} and {high 32 bits, tag}, tag = the low 32 bits of the pass number.  An aligned 8-byte store is
// indivisible on NVLink, on PCIe and in host memory, so a word either still carries an old tag or is complete: the
// receiver polls the words themselves and no flag, no fence and no second round trip are needed.  Tags never repeat
// within a buffer's reuse distance (two buffers alternate; 2^32 passes to wrap).
//
// Mailbox layout (per rank, device memory): data[2][kXrMaxRanks][kXrSlotDoubles] of such word pairs (16 bytes per value).
// Two buffers alternate with the pass number: a rank can only start pass s + 1 after it has read every slot of pass s,
// and nobody can finish pass s + 1 (and go on to overwrite buffer s % 2 with pass s + 2) before every rank has sent its
// vector of pass s + 1, i.e. finished reading pass s.  The host buffer uses the same word pairs.
#pragma once

#include "device_utils.cuh"
#include "kernels.hpp"

namespace dq4ml {

__device__ __forceinline__ unsigned long long* xr_slot(unsigned char* box, int buf, int r) {
  return reinterpret_cast<unsigned long long*>(box) + (static_cast<size_t>(buf) * kXrMaxRanks + r) * (2 * kXrSlotDoubles);
}
// plain (weak) 8-byte stores: each word is self-validating, ordering between words does not matter
__device__ __forceinline__ void ll_store(unsigned long long* pair, double v, unsigned tag) {
  const unsigned long long bits = static_cast<unsigned long long>(__double_as_longlong(v));
  const unsigned long long t = static_cast<unsigned long long>(tag) << 32;
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(pair), "l"((bits & 0xffffffffull) | t), "l"((bits >> 32) | t) : "memory");
}
__device__ __forceinline__ bool ll_load(const unsigned long long* pair, unsigned tag, double* v) {
  unsigned long long a, b;
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(pair) : "memory");
  if (static_cast<unsigned>(a >> 32) != tag || static_cast<unsigned>(b >> 32) != tag) return false;
  *v = __longlong_as_double(static_cast<long long>((a & 0xffffffffull) | (b << 32)));
  return true;
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
  const unsigned host_tag = static_cast<unsigned>(ep.host_seq);
  if (ep.world <= 1) {
    for (int k = tid; k < len; k += nthr) ll_store(ep.host_out + 2 * k, result[k], host_tag);
    return;
  }
  __shared__ int s_timeout;
  const int buf = static_cast<int>(ep.xr_seq & 1ull);
  const unsigned tag = static_cast<unsigned>(ep.xr_seq);
  if (tid == 0) s_timeout = 0;
  // 1. this rank's vector into its slot of every rank's mailbox (plain stores over NVLink; own mailbox included)
  for (int e = tid; e < ep.world * len; e += nthr) {
    const int r = e / len, k = e - r * len;
    ll_store(xr_slot(ep.peer[r], buf, ep.rank) + 2 * k, result[k], tag);
  }
  __syncthreads();  // (s_timeout initialised)
  // 2. collect from the own mailbox: a word is polled until it carries this pass's tag
  unsigned char* own = ep.peer[ep.rank];
  const unsigned long long t0 = global_timer_ns();
  auto fetch = [&](int r, int k, double* v) {
    const unsigned long long* src = xr_slot(own, buf, r) + 2 * k;
    unsigned spins = 0;
    while (!ll_load(src, tag, v)) {
      if ((++spins & 1023u) == 0u && global_timer_ns() - t0 > 30ull * 1000000000ull) {  // a rank died: give up, do not hang the GPU
        s_timeout = 1;
        return false;
      }
    }
    return true;
  };
  // 3. fold in rank order (bit-identical on every rank) or hand the host every rank's vector, word by word
  if (ep.gather != 0) {
    for (int e = tid; e < ep.world * len; e += nthr) {
      double v;
      if (fetch(e / len, e % len, &v)) ll_store(ep.host_out + 2 * e, v, host_tag);
    }
  } else {
    for (int k = tid; k < len; k += nthr) {
      double v = 0.0;
      bool ok = true;
      if (k < nacc) {
        const int kind = kind_of(k);
        ok = fetch(0, k, &v);
        for (int r = 1; r < ep.world && ok; ++r) {
          double o;
          ok = fetch(r, k, &o);
          v = combine(kind, v, o);
        }
      } else {
        v = result[k];  // constants are this rank's own
      }
      if (ok) ll_store(ep.host_out + 2 * k, v, host_tag);
    }
  }
  __syncthreads();
  if (s_timeout != 0 && tid == 0) ll_store(ep.host_out, 0.0, kXrErrorTag);
}

}  // namespace dq4ml
