// filter_kernels.cu -- the DQ section of the lab (DataQuality4MachineLearningApp.java:68-90: rule 1 -> WHERE > 0 ->
// rule 2 -> WHERE > 0, projecting (guest, price)) as ONE stable compaction pass, persistent-CTA version.
//
// Shape (one CTA per SM, 544 threads, ~200 KB of shared memory):
//   * a producer warp takes tile tickets (atomicAdd, so tiles are numbered in the order running CTAs ask for them -- no
//     assumption about the order in which the hardware dispatches CTAs) and fills a 4-stage shared-memory ring with TMA
//     bulk copies (cp.async.bulk global -> shared, completion on an mbarrier): 32 KB of prices + 16 KB of guests per
//     4096-row tile, ~100 KB in flight per SM without holding a single register;
//   * 16 consumer warps take a landed tile into registers: lane l of warp w owns rows l, l + 32, ... of the warp's
//     256-row slice, so one __ballot_sync per 32 rows gives every survivor its position; the tile's survivor count is
//     published to tile_state[tile] and the survivors are compacted in place in the stage;
//   * the write-out of a tile happens one tile LATER than its count: a CTA publishes the count of tile k right after
//     the iteration's first barrier and only then turns to the offset of tile k-1, so every count an offset needs has
//     been public for about a tile time and the look-back does not wait.  And it needs counts only, never another CTA's
//     prefix: a CTA knows the inclusive prefix of its own previous tile, so
//         offset(t) = prefix(previous tile of this CTA) + sum of the counts of the tiles in between
//     ~147 status words, one per thread, requested at the top of the iteration (their L2 round trip overlaps the wait for
//     the next tile) and summed across the CTA after the count has been published.  The chain of look-backs that paces
//     a classic decoupled look-back scan does not exist here;
//   * coalesced stores of the compacted stage; each warp then hands its share of the stage back to the producer through
//     the stage's mbarrier (16 arrivals) instead of a third CTA-wide barrier.
// Data is read once and written once (12 B per row in, 12 B per survivor out), order preserved.
//
// Variants measured on the way (profiles/r2_filter_variants.jsonl; fraction of the measured HBM peak at 10^9 rows, boxes
// differ by ~8 %): three CTA-wide barriers per tile, look-back loads where they are summed 0.72-0.79; + early request
// of the look-back words 0.76; + per-warp stage release 0.77; both (shipped) 0.80; the producer requesting its next
// ticket ahead of time 0.69-0.79 (tickets held back delay everybody who needs their count); survivors stored straight
// from registers, no parking 0.76 (0.68 at 4 x 10^9: each warp store is a contiguous but sector-unaligned span).  Two
// instructive failures: the look-back moved BEFORE the barrier and the count's publication 0.40 (every CTA then waits
// for every other CTA's previous publication before it publishes itself: a grid-wide barrier per tile at L2 latency);
// every warp fetching all ~147 status words for itself 0.01 (the status words sit on a handful of L2 lines that all
// 148 CTAs read and write, and 16x the strong loads on those lines starve the very stores they wait for).
//
// The round-1 kernel (one short-lived CTA per tile, survivors parked while the whole CTA walks back over its
// predecessors' words) measured 0.64; it stays in dq_kernels.cu as the path for columns that are not 16-byte aligned
// (bulk copies need that).
#include <cstdlib>

#include "device_utils.cuh"
#include "kernels.hpp"

namespace dq4ml {
namespace {

constexpr int kPfConsumers = 512;
constexpr int kPfWarps = kPfConsumers / 32;
constexpr int kPfThreads = kPfConsumers + 32;  // + the producer warp
constexpr unsigned long long kPfFlag = 1ull << 63;
constexpr int kPfDefaultVariant = 3;  // the measured-best combination of the kernel's V switches (profiles/r2_filter_variants.jsonl)

// PER_LANE rows per consumer thread and tile, STAGES ring slots: <8, 4> = 4096-row tiles of 48 KB (shipped).  Measured
// and dropped: <6, 6> = 3072-row tiles (0.68 against 0.72 on the same box) and 2048-row tiles.
template <int PER_LANE, int STAGES>
struct PfShape {
  static constexpr int kTile = PER_LANE * kPfConsumers;
  static constexpr int kWarpRows = 32 * PER_LANE;
  static constexpr int kPriceBytes = kTile * 8, kGuestBytes = kTile * 4, kStageBytes = kPriceBytes + kGuestBytes;
  struct Smem {
    unsigned char stage[STAGES][kStageBytes];  // raw rows of a tile, then (in place) its compacted survivors
    unsigned long long full[STAGES];
    unsigned long long empty[STAGES];
    long long tile[STAGES];
    unsigned int tot[2][kPfWarps];
    unsigned int rule1[2][kPfWarps];
    unsigned long long part[kPfWarps];
  };
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
// V: bit 0 = the look-back words are requested at the top of the iteration instead of where they are summed; bit 1 = each
// warp hands the stage back through the mbarrier (no third CTA-wide barrier); bit 2 = the producer requests its next
// ticket while the current tile is being issued.
template <int PER_LANE, int STAGES, int V>
__global__ void __launch_bounds__(kPfThreads, 1)
    dq_filter_stream_kernel(const int* __restrict__ guest, const double* __restrict__ price, const int64_t n,
                            const int64_t n_tiles, const DqThresholds t, int* __restrict__ guest_out,
                            double* __restrict__ price_out, unsigned long long* __restrict__ tile_state,
                            unsigned long long* __restrict__ counters) {
  using Shape = PfShape<PER_LANE, STAGES>;
  extern __shared__ __align__(128) unsigned char pf_raw[];
  typename Shape::Smem& sm = *reinterpret_cast<typename Shape::Smem*>(pf_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&sm.full[s]), 1);
      mbar_init(smem_u32(&sm.empty[s]), (V & 2) ? kPfWarps : 1);  // per-warp hand-back, or one arrival after a barrier
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kPfWarps) {
    // ------------------------------------------------------------------------------------------------ producer warp
    long long next_ticket = 0;
    if ((V & 4) && lane == 0) next_ticket = static_cast<long long>(atomicAdd(&counters[2], 1ull));
    for (int it = 0;; ++it) {
      const int s = it % STAGES;
      long long ticket = 0;
      if (lane == 0) {
        if (it >= STAGES) mbar_wait(smem_u32(&sm.empty[s]), ((it / STAGES) - 1) & 1);
        ticket = (V & 4) ? next_ticket : static_cast<long long>(atomicAdd(&counters[2], 1ull));
        sm.tile[s] = ticket < n_tiles ? ticket : -1;
      }
      ticket = __shfl_sync(0xffffffffu, ticket, 0);
      const uint32_t bar = smem_u32(&sm.full[s]);
      if (ticket >= n_tiles) {
        if (lane == 0) mbar_arrive(bar);
        break;
      }
      const int64_t row0 = ticket * Shape::kTile;
      double* s_price = reinterpret_cast<double*>(sm.stage[s]);
      int* s_guest = reinterpret_cast<int*>(sm.stage[s] + Shape::kPriceBytes);
      if (row0 + Shape::kTile <= n) {
        if (lane == 0) {
          mbar_arrive_expect_tx(bar, Shape::kStageBytes);
          bulk_g2s(smem_u32(s_price), price + row0, Shape::kPriceBytes, bar);
          bulk_g2s(smem_u32(s_guest), guest + row0, Shape::kGuestBytes, bar);
        }
      } else {
        // the ragged last tile: plain loads by the warp; rows past the end become price 0, which rule 1 drops
        for (int i = lane; i < Shape::kTile; i += 32) {
          const bool in = row0 + i < n;
          s_guest[i] = in ? guest[row0 + i] : 0;
          s_price[i] = in ? price[row0 + i] : 0.0;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar);
      }
      if ((V & 4) && lane == 0) next_ticket = static_cast<long long>(atomicAdd(&counters[2], 1ull));  // in flight during the next wait
    }
    return;
  }

  // -------------------------------------------------------------------------------------------------- consumer warps
  long long prev_tile = -1;            // last tile this CTA wrote out
  unsigned long long prev_incl = 0;    // survivors in tiles [0, prev_tile]
  unsigned long long rule1_total = 0;  // thread 0 only
  long long pend_tile = -1;            // tile counted and parked (in stage pend_stage), not yet written out
  int pend_stage = 0;
  unsigned pend_agg = 0;
  const unsigned lt = (1u << lane) - 1u;

  for (int it = 0;; ++it) {
    const int s = it % STAGES, buf = it & 1;
    // (0) the status words the pending tile's offset needs were published by the other CTAs about a tile time ago (a CTA
    // publishes a tile's count right after the barrier below and turns to the previous tile's offset only afterwards).
    // Thread t asks for the word of tile pend - 1 - t: ~147 threads, five warp-wide requests per CTA and tile -- the
    // status words sit on a handful of L2 lines that every CTA reads and writes, and those lines tolerate no more.
    const long long first_idx = pend_tile - 1 - tid;
    const bool have_first = pend_tile >= 0 && first_idx > prev_tile;
    unsigned long long first_w = 0;
    if ((V & 1) && have_first) first_w = ld_state(&tile_state[first_idx]);

    // (1) the next tile: rows into registers, one ballot per 32 rows
    mbar_wait(smem_u32(&sm.full[s]), (it / STAGES) & 1);
    const long long tile = sm.tile[s];
    double* s_price = reinterpret_cast<double*>(sm.stage[s]);
    int* s_guest = reinterpret_cast<int*>(sm.stage[s] + Shape::kPriceBytes);
    int g[PER_LANE];
    double p[PER_LANE];
    unsigned mask[PER_LANE];
    if (tile >= 0) {
      const int base = warp * Shape::kWarpRows + lane;
#pragma unroll
      for (int j = 0; j < PER_LANE; ++j) {
        g[j] = s_guest[base + 32 * j];
        p[j] = s_price[base + 32 * j];
      }
      unsigned warp_total = 0, only_rule2 = 0;
#pragma unroll
      for (int j = 0; j < PER_LANE; ++j) {
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
    }
    consumer_sync();  // warp totals visible; every thread holds its rows in registers, so the stage may be overwritten
    // (2) publish the new tile's count FIRST (others need it about one tile time from now), then compact in place
    unsigned aggregate = 0;
    if (tile >= 0) {
      unsigned warp_base = 0;
#pragma unroll
      for (int w = 0; w < kPfWarps; ++w) {
        const unsigned v = sm.tot[buf][w];
        if (w < warp) warp_base += v;
        aggregate += v;
      }
      if (tid == 0) {
        st_state(&tile_state[tile], kPfFlag | aggregate);
#pragma unroll
        for (int w = 0; w < kPfWarps; ++w) rule1_total += sm.rule1[buf][w];
      }
      unsigned run = warp_base;
#pragma unroll
      for (int j = 0; j < PER_LANE; ++j) {  // survivors of one ballot land on consecutive words
        if ((mask[j] >> lane) & 1u) {
          const unsigned at = run + __popc(mask[j] & lt);
          s_guest[at] = g[j];
          s_price[at] = p[j];
        }
        run += __popc(mask[j]);
      }
    }
    // (3) the pending tile: offset = prev_incl + counts of the tiles strictly between prev_tile and pend_tile, then
    // coalesced stores of its survivors (parked one iteration ago, visible since the barrier above)
    if (pend_tile >= 0) {
      unsigned long long sum = 0;
      for (long long idx = first_idx; idx > prev_tile; idx -= kPfConsumers) {
        unsigned long long w = ((V & 1) && idx == first_idx) ? first_w : ld_state(&tile_state[idx]);
        while ((w & kPfFlag) == 0ull) {
          __nanosleep(100);  // a straggler: polling would keep its store from getting through to the same L2 line
          w = ld_state(&tile_state[idx]);
        }
        sum += w & ~kPfFlag;
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, off);
      if (lane == 0) sm.part[warp] = sum;
      consumer_sync();
      unsigned long long exclusive = prev_incl;
#pragma unroll
      for (int w = 0; w < kPfWarps; ++w) exclusive += sm.part[w];
      const double* k_price = reinterpret_cast<const double*>(sm.stage[pend_stage]);
      const int* k_guest = reinterpret_cast<const int*>(sm.stage[pend_stage] + Shape::kPriceBytes);
      for (unsigned i = tid; i < pend_agg; i += kPfConsumers) {
        guest_out[exclusive + i] = k_guest[i];
        price_out[exclusive + i] = k_price[i];
      }
      prev_tile = pend_tile;
      prev_incl = exclusive + pend_agg;
      if (tid == 0 && pend_tile == n_tiles - 1) counters[0] = prev_incl;
      if (V & 2) {
        // this warp is done with the stage (its reads above, and sm.part); its parking writes of an iteration ago are
        // generic-proxy accesses that the stage's next bulk copy (async proxy) must be ordered after
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&sm.empty[pend_stage]));
      } else {
        consumer_sync();  // every read of the stage (and of sm.part) is done
        if (tid == 0) {
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          mbar_arrive(smem_u32(&sm.empty[pend_stage]));
        }
      }
    }
    if (tile < 0) break;
    pend_tile = tile;
    pend_stage = s;
    pend_agg = aggregate;
  }
  if (tid == 0 && rule1_total != 0ull) atomicAdd(&counters[1], rule1_total);
}

template <int PER_LANE, int STAGES, int V>
cudaError_t launch_shape(const int32_t* guest, const double* price, int64_t n, const DqThresholds& t, int32_t* guest_out,
                         double* price_out, unsigned long long* tile_state, unsigned long long* counters, int sm_count,
                         cudaStream_t stream) {
  using Shape = PfShape<PER_LANE, STAGES>;
  auto kernel = dq_filter_stream_kernel<PER_LANE, STAGES, V>;
  static bool smem_opt_in[64] = {false};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !smem_opt_in[dev]) {
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sizeof(typename Shape::Smem)));
    if (e != cudaSuccess) return e;
    smem_opt_in[dev] = true;
  }
  const int64_t n_tiles = (n + Shape::kTile - 1) / Shape::kTile;
  int64_t grid = sm_count > 0 ? sm_count : 148;
  if (grid > n_tiles) grid = n_tiles;
  kernel<<<static_cast<unsigned>(grid), kPfThreads, sizeof(typename Shape::Smem), stream>>>(
      guest, price, n, n_tiles, t, guest_out, price_out, tile_state, counters);
  return cudaGetLastError();
}

}  // namespace

bool dq_filter_stream_usable(const int32_t* guest, const double* price) {
  return ((reinterpret_cast<uintptr_t>(guest) | reinterpret_cast<uintptr_t>(price)) & 15u) == 0;
}

// tile_state [0, ceil(n / kFilterStateTile)) and counters[0..2] must be zero (the caller's memsets, on `stream`).
cudaError_t launch_dq_filter_stream(const int32_t* guest, const double* price, int64_t n, const DqThresholds& t,
                                    int32_t* guest_out, double* price_out, unsigned long long* tile_state,
                                    unsigned long long* counters, int sm_count, cudaStream_t stream) {
  static const int variant = [] {
    const char* env = std::getenv("DQ4ML_FILTER_VARIANT");  // A/B switch, see the kernel's V parameter: 0 or 3
    return env != nullptr ? std::atoi(env) : kPfDefaultVariant;
  }();
#define DQ4ML_PF_LAUNCH(PL, ST, VV) \
  return launch_shape<PL, ST, VV>(guest, price, n, t, guest_out, price_out, tile_state, counters, sm_count, stream)
  if ((variant & 7) == 0) DQ4ML_PF_LAUNCH(8, 4, 0);
  DQ4ML_PF_LAUNCH(8, 4, kPfDefaultVariant);
#undef DQ4ML_PF_LAUNCH
}

}  // namespace dq4ml
