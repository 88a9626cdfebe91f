// filter_kernels.cu -- the DQ section of the lab (DataQuality4MachineLearningApp.java:68-90: rule 1 -> WHERE > 0 ->
// rule 2 -> WHERE > 0, projecting (guest, price)) as ONE stable compaction pass, persistent-CTA version.
//
// Shape (one CTA per SM, 544 threads):
//   * a producer warp takes tile tickets (atomicAdd, so tiles are numbered in the order running CTAs ask for them -- no
//     assumption about the order in which the hardware dispatches CTAs) and fills a 4-stage shared-memory ring with
//     TMA bulk copies (cp.async.bulk global -> shared, completion on an mbarrier): 32 KB of prices + 16 KB of guests per
//     4096-row tile, up to 144 KB in flight per SM without holding a single register;
//   * 16 consumer warps take a landed tile: lane l of warp w owns rows l, l + 32, ... of the warp's 256-row slice, so
//     one __ballot_sync per 32 rows gives every survivor its position; the tile's survivor count is published to
//     tile_state[tile] at once, and the survivors are compacted in place in the stage;
//   * the write-out of a tile happens one tile LATER than its count: by then every other CTA has had a whole tile time
//     to publish the counts this tile's offset needs, so the look-back does not wait.  And it needs counts only, never
//     another CTA's prefix: a CTA knows the inclusive prefix of its own previous tile, so
//         offset(t) = prefix(previous tile of this CTA) + sum of the counts of the tiles in between
//     which is one L2 round trip over ~147 status words.  The chain of look-backs that paces a classic decoupled
//     look-back scan does not exist here;
//   * coalesced stores of the compacted stage, then the stage goes back to the producer.
// Data is read once and written once (12 B per row in, 12 B per survivor out), order preserved.
//
// The round-1 kernel (one short-lived CTA per tile, survivors parked while the whole CTA walks back over its
// predecessors' words) measured 0.64 of the HBM roofline; it stays in dq_kernels.cu as the path for columns that are not
// 16-byte aligned (bulk copies need that).
#include "device_utils.cuh"
#include "kernels.hpp"

namespace dq4ml {
namespace {

constexpr int kPfConsumers = 512;
constexpr int kPfWarps = kPfConsumers / 32;
constexpr int kPfThreads = kPfConsumers + 32;  // + the producer warp
constexpr int kPfTile = 4096;
constexpr int kPfPerLane = kPfTile / kPfConsumers;  // 8 rows per consumer thread
constexpr int kPfWarpRows = 32 * kPfPerLane;        // 256
constexpr int kPfStages = 4;
constexpr int kPfPriceBytes = kPfTile * 8, kPfGuestBytes = kPfTile * 4;
constexpr int kPfStageBytes = kPfPriceBytes + kPfGuestBytes;  // 48 KB
constexpr unsigned long long kPfFlag = 1ull << 63;

struct PfSmem {
  unsigned char stage[kPfStages][kPfStageBytes];
  unsigned long long full[kPfStages];
  unsigned long long empty[kPfStages];
  long long tile[kPfStages];
  unsigned int tot[2][kPfWarps];
  unsigned int rule1[2][kPfWarps];
  unsigned long long part[kPfWarps];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  } while (!done);
}
// TMA bulk copy global -> this CTA's shared memory; both addresses 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kPfConsumers) : "memory"); }

__device__ __forceinline__ unsigned long long ld_state(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_state(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// counters: [0] survivors, [1] rows after rule 1, [2] tile ticket -- all zero at launch, like tile_state[0, n_tiles).
__global__ void __launch_bounds__(kPfThreads, 1)
    dq_filter_stream_kernel(const int* __restrict__ guest, const double* __restrict__ price, const int64_t n,
                            const int64_t n_tiles, const DqThresholds t, int* __restrict__ guest_out,
                            double* __restrict__ price_out, unsigned long long* __restrict__ tile_state,
                            unsigned long long* __restrict__ counters) {
  extern __shared__ __align__(128) unsigned char pf_raw[];
  PfSmem& sm = *reinterpret_cast<PfSmem*>(pf_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int s = 0; s < kPfStages; ++s) {
      mbar_init(smem_u32(&sm.full[s]), 1);
      mbar_init(smem_u32(&sm.empty[s]), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kPfWarps) {
    // ------------------------------------------------------------------------------------------------ producer warp
    for (int it = 0;; ++it) {
      const int s = it % kPfStages;
      long long ticket = 0;
      if (lane == 0) {
        if (it >= kPfStages) mbar_wait(smem_u32(&sm.empty[s]), ((it / kPfStages) - 1) & 1);
        ticket = static_cast<long long>(atomicAdd(&counters[2], 1ull));
        sm.tile[s] = ticket < n_tiles ? ticket : -1;
      }
      ticket = __shfl_sync(0xffffffffu, ticket, 0);
      const uint32_t bar = smem_u32(&sm.full[s]);
      if (ticket >= n_tiles) {
        if (lane == 0) mbar_arrive(bar);
        break;
      }
      const int64_t row0 = ticket * kPfTile;
      double* s_price = reinterpret_cast<double*>(sm.stage[s]);
      int* s_guest = reinterpret_cast<int*>(sm.stage[s] + kPfPriceBytes);
      if (row0 + kPfTile <= n) {
        if (lane == 0) {
          mbar_arrive_expect_tx(bar, kPfStageBytes);
          bulk_g2s(smem_u32(s_price), price + row0, kPfPriceBytes, bar);
          bulk_g2s(smem_u32(s_guest), guest + row0, kPfGuestBytes, bar);
        }
      } else {
        // the ragged last tile: plain loads by the warp; rows past the end become price 0, which rule 1 drops
        for (int i = lane; i < kPfTile; i += 32) {
          const bool in = row0 + i < n;
          s_guest[i] = in ? guest[row0 + i] : 0;
          s_price[i] = in ? price[row0 + i] : 0.0;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar);
      }
    }
    return;
  }

  // -------------------------------------------------------------------------------------------------- consumer warps
  long long prev_tile = -1;             // last tile this CTA wrote out
  unsigned long long prev_incl = 0;     // survivors in tiles [0, prev_tile]
  unsigned long long rule1_total = 0;   // thread 0 only
  long long pend_tile = -1;             // tile counted and parked, not yet written out
  int pend_stage = 0;
  unsigned pend_agg = 0;

  auto write_out = [&]() {
    // offset(pend_tile) = prev_incl + counts of the tiles strictly between prev_tile and pend_tile
    unsigned long long sum = 0;
    for (long long idx = pend_tile - 1 - tid; idx > prev_tile; idx -= kPfConsumers) {
      unsigned long long w;
      do { w = ld_state(&tile_state[idx]); } while ((w & kPfFlag) == 0ull);
      sum += w & ~kPfFlag;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, off);
    if (lane == 0) sm.part[warp] = sum;
    consumer_sync();
    unsigned long long exclusive = prev_incl;
#pragma unroll
    for (int w = 0; w < kPfWarps; ++w) exclusive += sm.part[w];
    const double* s_price = reinterpret_cast<const double*>(sm.stage[pend_stage]);
    const int* s_guest = reinterpret_cast<const int*>(sm.stage[pend_stage] + kPfPriceBytes);
    for (unsigned i = tid; i < pend_agg; i += kPfConsumers) {
      guest_out[exclusive + i] = s_guest[i];
      price_out[exclusive + i] = s_price[i];
    }
    prev_tile = pend_tile;
    prev_incl = exclusive + pend_agg;
    if (tid == 0 && pend_tile == n_tiles - 1) counters[0] = prev_incl;
    consumer_sync();  // every read of the stage (and of sm.part) is done
    if (tid == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy accesses before the next bulk copy
      mbar_arrive(smem_u32(&sm.empty[pend_stage]));
    }
  };

  for (int it = 0;; ++it) {
    const int s = it % kPfStages, buf = it & 1;
    mbar_wait(smem_u32(&sm.full[s]), (it / kPfStages) & 1);
    const long long tile = sm.tile[s];
    if (tile < 0) break;
    double* s_price = reinterpret_cast<double*>(sm.stage[s]);
    int* s_guest = reinterpret_cast<int*>(sm.stage[s] + kPfPriceBytes);
    const int base = warp * kPfWarpRows + lane;
    int g[kPfPerLane];
    double p[kPfPerLane];
#pragma unroll
    for (int j = 0; j < kPfPerLane; ++j) {
      g[j] = s_guest[base + 32 * j];
      p[j] = s_price[base + 32 * j];
    }
    unsigned mask[kPfPerLane];
    unsigned warp_total = 0, only_rule2 = 0;
#pragma unroll
    for (int j = 0; j < kPfPerLane; ++j) {
      const bool k1 = dq_keep_rule1(p[j], t);
      const bool k = k1 && !(g[j] < t.min_guest && p[j] > t.max_price);
      mask[j] = __ballot_sync(0xffffffffu, k);
      only_rule2 += (k1 && !k) ? 1u : 0u;
      warp_total += __popc(mask[j]);
    }
    const unsigned rule1 = warp_total + __reduce_add_sync(0xffffffffu, only_rule2);
    if (lane == 0) {
      sm.tot[buf][warp] = warp_total;
      sm.rule1[buf][warp] = rule1;
    }
    consumer_sync();  // warp totals visible; every thread has its rows in registers, so the stage may be overwritten
    unsigned warp_base = 0, aggregate = 0;
#pragma unroll
    for (int w = 0; w < kPfWarps; ++w) {
      const unsigned v = sm.tot[buf][w];
      if (w < warp) warp_base += v;
      aggregate += v;
    }
    if (tid == 0) {
      st_state(&tile_state[tile], kPfFlag | aggregate);  // published a whole tile time before anyone needs it
#pragma unroll
      for (int w = 0; w < kPfWarps; ++w) rule1_total += sm.rule1[buf][w];
    }
    {  // compact in place: survivors of one ballot land on consecutive words
      const unsigned lt = (1u << lane) - 1u;
      unsigned run = warp_base;
#pragma unroll
      for (int j = 0; j < kPfPerLane; ++j) {
        if ((mask[j] >> lane) & 1u) {
          const unsigned local = run + __popc(mask[j] & lt);
          s_guest[local] = g[j];
          s_price[local] = p[j];
        }
        run += __popc(mask[j]);
      }
    }
    if (pend_tile >= 0) write_out();
    pend_tile = tile;
    pend_stage = s;
    pend_agg = aggregate;
  }
  if (pend_tile >= 0) write_out();
  if (tid == 0 && rule1_total != 0ull) atomicAdd(&counters[1], rule1_total);
}

}  // namespace

bool dq_filter_stream_usable(const int32_t* guest, const double* price) {
  return ((reinterpret_cast<uintptr_t>(guest) | reinterpret_cast<uintptr_t>(price)) & 15u) == 0;
}

// tile_state [0, ceil(n / 4096)) and counters[0..2] must be zero (the caller's memsets, on `stream`).
cudaError_t launch_dq_filter_stream(const int32_t* guest, const double* price, int64_t n, const DqThresholds& t,
                                    int32_t* guest_out, double* price_out, unsigned long long* tile_state,
                                    unsigned long long* counters, int sm_count, cudaStream_t stream) {
  static bool smem_opt_in[64] = {false};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !smem_opt_in[dev]) {
    e = cudaFuncSetAttribute(dq_filter_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sizeof(PfSmem)));
    if (e != cudaSuccess) return e;
    smem_opt_in[dev] = true;
  }
  const int64_t n_tiles = (n + kPfTile - 1) / kPfTile;
  int64_t grid = sm_count > 0 ? sm_count : 148;
  if (grid > n_tiles) grid = n_tiles;
  dq_filter_stream_kernel<<<static_cast<unsigned>(grid), kPfThreads, sizeof(PfSmem), stream>>>(
      guest, price, n, n_tiles, t, guest_out, price_out, tile_state, counters);
  return cudaGetLastError();
}

}  // namespace dq4ml
