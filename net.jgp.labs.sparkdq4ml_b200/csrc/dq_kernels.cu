// dq_kernels.cu -- the DQ section of the lab (DataQuality4MachineLearningApp.java:68-90) as sm_100a kernels:
// the two UDF columns, and rule1 -> filter -> rule2 -> filter fused into one stable compaction pass.
// Plus the synthetic-data generator and the row digest used by the benchmarks/tests.
#include <cstdlib>

#include "device_utils.cuh"
#include "kernels.hpp"

namespace dq4ml {

// filter_kernels.cu
bool dq_filter_stream_usable(const int32_t* guest, const double* price);
cudaError_t launch_dq_filter_stream(const int32_t* guest, const double* price, int64_t n, const DqThresholds& t,
                                    int32_t* guest_out, double* price_out, unsigned long long* tile_state,
                                    unsigned long long* counters, int sm_count, cudaStream_t stream);

namespace {

__host__ DqThresholds thresholds(const dq4ml_dq_params& p) {
  DqThresholds t;
  t.min_price = p.min_price;
  t.max_price = p.max_price;
  t.min_guest = p.min_guest;
  return t;
}

// ------------------------------------------------------------------------------------------ UDF columns
// minimumPriceRule over a column: MinimumPriceDataQualityUdf.call -> checkMinimumPrice
// (MinimumPriceDataQualityService.java:7-13).  A null (valid == 0) is the reference's NullPointerException.
__global__ void __launch_bounds__(256) udf_min_price_kernel(const double* __restrict__ price,
                                                            const uint8_t* __restrict__ valid, int64_t n,
                                                            double* __restrict__ out, double min_price,
                                                            int* __restrict__ err_flag) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 2;
  for (int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 2; i < n; i += stride) {
    if (i + 2 <= n) {
      const double2 p = ldg_stream_d2(price + i);
      if (valid != nullptr && (valid[i] == 0 || valid[i + 1] == 0)) *err_flag = 1;
      stg_stream_d2(out + i, make_double2(p.x < min_price ? -1.0 : p.x, p.y < min_price ? -1.0 : p.y));
    } else {
      if (valid != nullptr && valid[i] == 0) *err_flag = 1;
      const double p = price[i];
      out[i] = p < min_price ? -1.0 : p;
    }
  }
}

// priceCorrelationRule over columns: PriceCorrelationDataQualityUdf.call (null -> -1.0, .java:12-14) ->
// checkPriceRange (PriceCorrelationDataQualityService.java:5-10).
__global__ void __launch_bounds__(256) udf_price_corr_kernel(const double* __restrict__ price,
                                                             const int* __restrict__ guest,
                                                             const uint8_t* __restrict__ pvalid,
                                                             const uint8_t* __restrict__ gvalid, int64_t n,
                                                             double* __restrict__ out, double max_price,
                                                             int min_guest) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 2;
  for (int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 2; i < n; i += stride) {
    if (i + 2 <= n) {
      const double2 p = ldg_stream_d2(price + i);
      const int2 g = ldg_stream_i2(guest + i);
      const bool null0 = (pvalid != nullptr && pvalid[i] == 0) || (gvalid != nullptr && gvalid[i] == 0);
      const bool null1 = (pvalid != nullptr && pvalid[i + 1] == 0) || (gvalid != nullptr && gvalid[i + 1] == 0);
      double2 o;
      o.x = (null0 || (g.x < min_guest && p.x > max_price)) ? -1.0 : p.x;
      o.y = (null1 || (g.y < min_guest && p.y > max_price)) ? -1.0 : p.y;
      stg_stream_d2(out + i, o);
    } else {
      const bool nul = (pvalid != nullptr && pvalid[i] == 0) || (gvalid != nullptr && gvalid[i] == 0);
      const double p = price[i];
      out[i] = (nul || (guest[i] < min_guest && p > max_price)) ? -1.0 : p;
    }
  }
}

// ------------------------------------------------------------------------------------------ filter + compaction
// One CTA of 256 threads per tile of 4096 consecutive rows; warp w owns rows [512 w, 512 w + 512) of the tile and its
// lane l the rows l, l + 32, ... of that slice (coalesced 4- and 8-byte loads).  Tile t publishes its
// survivor count in tile_state[t] as {flag:2 | value:62}; flag 1 = "aggregate of this tile", flag 2 = "inclusive prefix
// up to this tile".  A tile finds its output offset by walking back over its predecessors' words (decoupled
// look-back), so the data is read once and written once, and the output order is the input order (Spark's filter is
// order preserving within a partition).  Survivors are first compacted in shared memory, then written out coalesced.
//
// Round-1 history (profiles/): 2048-row tiles with direct per-thread stores ran at 0.33 of the HBM roofline with 62 %
// of the warp stalls on the barrier behind warp 0's look-back -- with ~600 tiles in flight the walk back to the nearest
// published prefix took ~10 L2 round trips.  Doubling the tile and halving the CTAs in flight shortens that walk, and
// the other warps park their survivors in shared memory while warp 0 walks.
constexpr int kFilterThreads = 256;
constexpr int64_t kFilterTile = 4096;  // rows per tile: 16 per thread
constexpr size_t kFilterSmem = (kFilterTile + 4) * (sizeof(double) + sizeof(int));  // 48 KB of staged survivors
constexpr unsigned long long kFlagAgg = 1ull << 62, kFlagPrefix = 2ull << 62, kValueMask = (1ull << 62) - 1;

__device__ __forceinline__ unsigned long long ld_state(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_state(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// One short-lived CTA per tile; blockIdx order is dispatch order, so every predecessor of a running tile has started.
// (Two consecutive tiles per CTA with both tiles' loads issued up front -- to hide the second tile's DRAM latency behind
// the first tile's look-back -- is 70x slower: tile 2b+1 publishes its count only after tile 2b is completely done,
// which chains every CTA behind its predecessor.  Publishing both counts first removes the chain but also the overlap.)
// Measured alternatives (round 1, 10^9 rows): 512 x 4-row threads with per-thread stores 0.33 of the HBM roofline;
// 256 threads x 16 rows (128-bit loads, shuffle scans, shared-memory staging) 0.51; persistent CTAs with cp.async
// prefetch of the next tile and a ticket counter 0.20 (every round turns into a grid-wide barrier on the slowest CTA);
// 512 threads x 8 rows 0.43; the shipped striped shape with 1024 threads x 16 rows (one 192 KB CTA per SM, a single
// look-back step) 0.64 against 0.66 for 256 x 16 at four CTAs per SM -- the shape no longer matters.  Where the
// remaining third goes (10^9 rows, tools/filter_ceiling.cu and ablations of this kernel): a plain copy with the same
// loads, shared-memory parking and write-out takes 3.46 ms (6.9 TB/s); this kernel 5.61 ms; without the look-back
// 4.17 ms; without look-back and write-out 2.06 ms.  So the look-back costs 1.4 ms and the tile-local logic 0.7 ms.
// Reading three look-back windows in one round trip instead of one per step made it slower (6.09 ms, spills): the
// time is not in the number of steps but in waiting for the slowest of the ~600 tiles in flight to publish its count,
// and four CTAs per SM (shared memory and registers both full) are not enough other work to cover that.  An ablation of the 0.51 shape without look-back and write-out still took 4.15 ms per 10^9
// rows: the tile-local work was instruction-bound (~67 thread instructions per row: 5-step shuffle scans and 4-way
// bank-conflicted parking).  Hence the striped row assignment below: lane l of a warp owns rows l, l + 32, ... of the
// warp's 512-row slice, so one __ballot_sync per 32 rows replaces the scans and survivors of one ballot land on
// consecutive shared-memory words.
__global__ void __launch_bounds__(kFilterThreads, 4)
    dq_filter_compact_kernel(const int* __restrict__ guest, const double* __restrict__ price, const int64_t n,
                             const DqThresholds t, int* __restrict__ guest_out, double* __restrict__ price_out,
                             unsigned long long* __restrict__ tile_state, unsigned long long* __restrict__ counters) {
  extern __shared__ __align__(16) unsigned char filter_smem[];
  double* s_price = reinterpret_cast<double*>(filter_smem);              // survivors of the tile, in row order (+ <= 3 slack)
  int* s_guest = reinterpret_cast<int*>(filter_smem + (kFilterTile + 4) * sizeof(double));
  constexpr int kWarps = kFilterThreads / 32;
  constexpr int kPerLane = kFilterTile / kFilterThreads;  // 16 rows per thread
  __shared__ unsigned int s_tot[kWarps];
  __shared__ unsigned int s_rule1[kWarps];
  __shared__ unsigned long long s_part[kWarps];
  __shared__ bool s_found[kWarps];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t n_tiles = gridDim.x;
  const int64_t tile = blockIdx.x;
  {
    const int64_t warp_row = tile * kFilterTile + static_cast<int64_t>(warp) * (32 * kPerLane) + lane;
    int g[kPerLane];
    double p[kPerLane];
    if (warp_row - lane + 32 * kPerLane <= n) {
#pragma unroll
      for (int j = 0; j < kPerLane; ++j) {
        g[j] = __ldg(guest + warp_row + 32 * j);
        p[j] = __ldg(price + warp_row + 32 * j);
      }
    } else {
#pragma unroll
      for (int j = 0; j < kPerLane; ++j) {
        const bool in = warp_row + 32 * j < n;
        g[j] = in ? guest[warp_row + 32 * j] : 0;
        p[j] = in ? price[warp_row + 32 * j] : 0.0;  // price 0 is dropped by rule 1
      }
    }
    unsigned mask[kPerLane];
    unsigned warp_total = 0, only_rule2 = 0;  // rows that pass rule 1 but fall to rule 2 (rare): counted per thread
    // rows past the end of the data were loaded as price 0, which rule 1 drops: no bounds test is needed here
#pragma unroll
    for (int j = 0; j < kPerLane; ++j) {
      const bool k1 = dq_keep_rule1(p[j], t);
      const bool k = k1 && !(g[j] < t.min_guest && p[j] > t.max_price);
      mask[j] = __ballot_sync(0xffffffffu, k);
      only_rule2 += (k1 && !k) ? 1u : 0u;
      warp_total += __popc(mask[j]);
    }
    const unsigned rule1 = warp_total + __reduce_add_sync(0xffffffffu, only_rule2);
    if (lane == 0) {
      s_tot[warp] = warp_total;
      s_rule1[warp] = rule1;
    }
    __syncthreads();

    // every warp sums the 8 warp totals itself
    unsigned warp_base = 0, aggregate = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const unsigned v = s_tot[w];
      if (w < warp) warp_base += v;
      aggregate += v;
    }
    if (tid == 0) {
      unsigned r1 = 0;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) r1 += s_rule1[w];
      if (r1 != 0u) atomicAdd(&counters[1], static_cast<unsigned long long>(r1));
      // publish this tile's survivor count at once: successors can add it while we are still looking for our offset
      st_state(&tile_state[tile], (tile == 0 ? kFlagPrefix : kFlagAgg) | aggregate);
    }
    // park the survivors at their tile-local position (before the look-back, whose latency this hides): the survivors
    // of one ballot are consecutive words
    {
      const unsigned lt = (1u << lane) - 1u;
      unsigned run = warp_base;
#pragma unroll
      for (int j = 0; j < kPerLane; ++j) {
        if ((mask[j] >> lane) & 1u) {
          const unsigned local = run + __popc(mask[j] & lt);
          s_guest[local] = g[j];
          s_price[local] = p[j];
        }
        run += __popc(mask[j]);
      }
    }

    // Decoupled look-back by the whole CTA: thread i reads the word of tile (tile - 1 - i - 256 k) in step k, so one
    // step covers 256 predecessors (one L2 round trip).  A warp adds the counts of its 32 tiles up to and including
    // the nearest one that already holds a full prefix; the warps' results are chained in tile order.
    unsigned long long exclusive = 0;
    if (tile > 0) {
      int64_t first_idx = tile - 1;
      for (;;) {
        const int64_t my = first_idx - tid;
        unsigned long long w = kFlagPrefix;  // tiles before 0 contribute an (empty) prefix
        if (my >= 0) {
          do { w = ld_state(&tile_state[my]); } while ((w >> 62) == 0ull);
        }
        const unsigned has_prefix = __ballot_sync(0xffffffffu, (w >> 62) == 2ull);
        const int first = has_prefix ? __ffs(has_prefix) - 1 : 31;
        unsigned long long contrib = lane <= first ? (w & kValueMask) : 0ull;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) contrib += __shfl_down_sync(0xffffffffu, contrib, off);
        if (lane == 0) {
          s_part[warp] = contrib;
          s_found[warp] = has_prefix != 0u;
        }
        __syncthreads();
        bool done = false;
#pragma unroll
        for (int wv = 0; wv < kWarps; ++wv) {
          if (!done) {
            exclusive += s_part[wv];
            done = s_found[wv];
          }
        }
        __syncthreads();  // s_part / s_found are rewritten in the next step
        if (done) break;
        first_idx -= kFilterThreads;
      }
      if (tid == 0) st_state(&tile_state[tile], kFlagPrefix | (exclusive + aggregate));
    } else {
      __syncthreads();  // the parked survivors must be visible before the write-out
    }
    if (tid == 0 && tile == n_tiles - 1) counters[0] = exclusive + aggregate;

    // coalesced write-out (128-bit stores on the output's 16-byte grid were tried: the phase mismatch with shared
    // memory costs more in bank conflicts than the wider stores save -- 0.62 against 0.64 of the roofline)
    for (unsigned i = tid; i < aggregate; i += kFilterThreads) {
      guest_out[exclusive + i] = s_guest[i];
      price_out[exclusive + i] = s_price[i];
    }
  }
}

// ------------------------------------------------------------------------------------------ synthetic data
// Counter-based generator, bit-identical to oracle/dq4ml_oracle.c:orc_synth_generate (integer hashing plus
// individually rounded fp64 add/mul: explicit __dadd_rn/__dmul_rn so that nothing is contracted into an FMA).
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
  z ^= z >> 27; z *= 0x94D049BB133111EBull;
  z ^= z >> 31;
  return z;
}
constexpr unsigned long long kPhi = 0x9E3779B97F4A7C15ull;
__device__ __forceinline__ unsigned long long row_key(unsigned long long seed, unsigned long long r) { return mix64(seed + (r + 1) * kPhi); }
__device__ __forceinline__ unsigned long long stream_of(unsigned long long key, unsigned long long k) { return mix64(key + k * kPhi); }
__device__ __forceinline__ double u01(unsigned long long h) { return __dmul_rn(static_cast<double>(h >> 11), 0x1.0p-53); }

__global__ void __launch_bounds__(256) synth_kernel(const dq4ml_synth_params sp, const double thr1, const double thr2,
                                                    const int64_t row0, const int64_t n, int* __restrict__ guest,
                                                    double* __restrict__ price, double* __restrict__ extra,
                                                    const int64_t extra_stride) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const unsigned long long r = static_cast<unsigned long long>(row0 + i);
    const unsigned long long key = row_key(sp.seed_data, r);
    const int g = 1 + static_cast<int>(((stream_of(key, 1) >> 32) * 35ull) >> 32);
    double s = __dadd_rn(u01(stream_of(key, 2)), u01(stream_of(key, 3)));
    s = __dadd_rn(s, u01(stream_of(key, 4)));
    s = __dadd_rn(s, u01(stream_of(key, 5)));
    const double noise = __dmul_rn(__dadd_rn(s, -2.0), 2.6);
    double pr = __dadd_rn(static_cast<double>(5 * g + 22), noise);
    for (int k = 1; k < sp.d; ++k) {
      const double xk = __dmul_rn(u01(stream_of(key, 8ull + k)), static_cast<double>(1 + (k & 3)));
      if (extra != nullptr) extra[static_cast<int64_t>(k - 1) * extra_stride + i] = xk;
      const double gamma = static_cast<double>((k * 7) % 5 - 2) * 0.5;
      pr = __dadd_rn(pr, __dmul_rn(gamma, xk));
    }
    const unsigned long long akey = row_key(sp.seed_anom, r);
    const double ua = u01(mix64(akey));
    const double ub = u01(stream_of(akey, 1));
    if (ua < thr1) pr = __dadd_rn(3.0, __dmul_rn(12.0, ub));
    else if (ua < thr2 && g < 14) pr = __dadd_rn(100.0, __dmul_rn(50.0, ub));
    guest[i] = g;
    price[i] = pr;
  }
}

// Order-sensitive digest: two 64-bit sums of per-row hashes keyed by the row's position (integer adds commute, so the
// reduction order does not matter and atomics are exact).
__global__ void __launch_bounds__(256) rows_digest_kernel(const int* __restrict__ guest, const double* __restrict__ price,
                                                          const int64_t n, const unsigned long long seed,
                                                          unsigned long long* __restrict__ out) {
  unsigned long long a = 0, b = 0;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const unsigned long long bits = static_cast<unsigned long long>(__double_as_longlong(price[i]));
    unsigned long long h = mix64((seed + (static_cast<unsigned long long>(i) + 1) * kPhi) ^ bits);
    h = mix64(h + static_cast<unsigned long long>(static_cast<unsigned int>(guest[i])) * 0xD1B54A32D192ED03ull);
    a += h;
    b += mix64(h ^ (static_cast<unsigned long long>(i) * 0x2545F4914F6CDD1Dull));
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    a += __shfl_down_sync(0xffffffffu, a, off);
    b += __shfl_down_sync(0xffffffffu, b, off);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&out[0], a);
    atomicAdd(&out[1], b);
  }
}

// mean of the first (up to 256) values of each column: provisional centres for the shifted-sum summarizer
__global__ void __launch_bounds__(256) column_heads_kernel(const DeviceCols cols, const double* __restrict__ label,
                                                           const int label_slot, const int64_t n,
                                                           double* __restrict__ centres) {
  __shared__ double s_red[256 / 32];
  const int64_t m = n < 256 ? n : 256;
  // slots [0, cols.n) = feature columns, [cols.n, label_slot) = unused (0), label_slot = label
  for (int c = 0; c <= label_slot; ++c) {
    double v = 0.0;
    const bool live = c < cols.n || (c == label_slot && label != nullptr);
    if (live && threadIdx.x < m) {
      if (c < cols.n) {
        v = cols.dtype[c] == DQ4ML_I32 ? static_cast<double>(static_cast<const int*>(cols.ptr[c])[threadIdx.x])
                                       : static_cast<const double*>(cols.ptr[c])[threadIdx.x];
      } else {
        v = label[threadIdx.x];
      }
      if (!(fabs(v) <= 1.79e308)) v = 0.0;  // NaN/Inf must not become the centre
    }
    v = block_reduce(v, kSum, s_red);
    if (threadIdx.x == 0) centres[c] = (live && m > 0) ? v / static_cast<double>(m) : 0.0;
  }
}

int grid_for(int64_t work_items, int threads, int cap_blocks) {
  int64_t b = (work_items + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > cap_blocks) b = cap_blocks;
  return static_cast<int>(b);
}

}  // namespace

cudaError_t launch_udf_min_price(const double* price, const uint8_t* valid, int64_t n, double* out,
                                 const dq4ml_dq_params& p, int* err_flag, cudaStream_t stream, int64_t* launches) {
  if (n <= 0) return cudaSuccess;
  udf_min_price_kernel<<<grid_for((n + 1) / 2, 256, 148 * 16), 256, 0, stream>>>(price, valid, n, out, p.min_price, err_flag);
  ++*launches;
  return cudaGetLastError();
}

cudaError_t launch_udf_price_corr(const double* price, const int32_t* guest, const uint8_t* pvalid,
                                  const uint8_t* gvalid, int64_t n, double* out, const dq4ml_dq_params& p,
                                  cudaStream_t stream, int64_t* launches) {
  if (n <= 0) return cudaSuccess;
  udf_price_corr_kernel<<<grid_for((n + 1) / 2, 256, 148 * 16), 256, 0, stream>>>(price, guest, pvalid, gvalid, n, out,
                                                                                   p.max_price, p.min_guest);
  ++*launches;
  return cudaGetLastError();
}

size_t dq_filter_state_bytes(int64_t n) {
  // one status word per tile of the kernel with the smaller tile (the persistent kernel, 2048 rows); the
  // one-CTA-per-tile kernel uses the first half of them
  const int64_t tiles = (n + kFilterStateTile - 1) / kFilterStateTile;
  return static_cast<size_t>(tiles > 0 ? tiles : 1) * sizeof(unsigned long long);
}

cudaError_t launch_dq_filter(const int32_t* guest, const double* price, int64_t n, const dq4ml_dq_params& p,
                             int32_t* guest_out, double* price_out, unsigned long long* tile_state,
                             unsigned long long* counters, cudaStream_t stream, int64_t* launches) {
  cudaError_t e = cudaMemsetAsync(counters, 0, 3 * sizeof(unsigned long long), stream);  // survivors, after rule 1, tile ticket
  if (e != cudaSuccess || n <= 0) return e;
  e = cudaMemsetAsync(tile_state, 0, dq_filter_state_bytes(n), stream);
  if (e != cudaSuccess) return e;
  const int64_t tiles = (n + kFilterTile - 1) / kFilterTile;
  static bool smem_opt_in[64] = {false};
  static int sm_counts[64] = {0};
  int dev = 0;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
  // the persistent TMA-ring kernel (filter_kernels.cu) whenever the columns allow bulk copies; DQ4ML_FILTER_LEGACY=1
  // forces the one-CTA-per-tile kernel below (kept for columns that are not 16-byte aligned, and for A/B timing)
  static const bool legacy = getenv("DQ4ML_FILTER_LEGACY") != nullptr;
  if (!legacy && dq_filter_stream_usable(guest, price)) {
    int sms = 148;
    if (dev >= 0 && dev < 64) {
      if (sm_counts[dev] == 0 && (e = cudaDeviceGetAttribute(&sm_counts[dev], cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
      sms = sm_counts[dev];
    }
    ++*launches;
    return launch_dq_filter_stream(guest, price, n, thresholds(p), guest_out, price_out, tile_state, counters, sms, stream);
  }
  if (dev >= 0 && dev < 64 && !smem_opt_in[dev]) {
    e = cudaFuncSetAttribute(dq_filter_compact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kFilterSmem));
    if (e != cudaSuccess) return e;
    smem_opt_in[dev] = true;
  }
  const int64_t grid = tiles;
  dq_filter_compact_kernel<<<static_cast<unsigned>(grid), kFilterThreads, kFilterSmem, stream>>>(
      guest, price, n, thresholds(p), guest_out, price_out, tile_state, counters);
  ++*launches;
  return cudaGetLastError();
}

cudaError_t launch_synth(const dq4ml_synth_params& p, int64_t row0, int64_t n, int32_t* guest, double* price,
                         double* extra, int64_t extra_stride, cudaStream_t stream, int64_t* launches) {
  if (n <= 0) return cudaSuccess;
  synth_kernel<<<grid_for(n, 256, 148 * 32), 256, 0, stream>>>(p, p.p_low, p.p_low + p.p_high, row0, n, guest, price,
                                                               extra, extra_stride);
  ++*launches;
  return cudaGetLastError();
}

cudaError_t launch_rows_digest(const int32_t* guest, const double* price, int64_t n, uint64_t seed,
                               unsigned long long* out2, cudaStream_t stream, int64_t* launches) {
  cudaError_t e = cudaMemsetAsync(out2, 0, 2 * sizeof(unsigned long long), stream);
  if (e != cudaSuccess || n <= 0) return e;
  rows_digest_kernel<<<grid_for(n, 256, 148 * 16), 256, 0, stream>>>(guest, price, n, seed, out2);
  ++*launches;
  return cudaGetLastError();
}

cudaError_t launch_column_heads(const DeviceCols& cols, const double* label, int label_slot, int64_t n,
                                const Scratch& scratch, cudaStream_t stream, int64_t* launches) {
  column_heads_kernel<<<1, 256, 0, stream>>>(cols, label, label_slot, n, scratch.consts);
  ++*launches;
  return cudaGetLastError();
}

}  // namespace dq4ml

// =====================================================================================================================
// dq4ml_dq_filter_noalloc -- the DQ filter for callers that must not let this library allocate.
//
// Who needs it: a host process whose GPU memory already belongs to somebody else's pool (a Spark executor running an
// RMM-style allocator that reserved the whole device, for instance), or a caller on a latency-critical path where
// the device-wide synchronisation hidden in cudaMalloc/cudaFree is not acceptable.  The regular entry point,
// dq4ml_dq_filter, needs a dq4ml_ctx (whose creation allocates scratch) and grows its tile-status buffer with
// cudaMalloc; this one takes no ctx and touches no allocator at all.
//
// Where the scratch lives instead: the look-back words of the compaction kernel (one 64-bit word per 2048-row tile)
// plus the two row counters sit in the file-scope device array below.  It is part of the library image, so every GPU
// that loads libdq4ml.so carries these ~256 KB until the process exits, whether or not this entry point is ever used.
// That fixed size is what bounds the call: at most DQ4ML_NOALLOC_MAX_ROWS rows per call.
#include <mutex>

#include "context.hpp"

namespace dq4ml {
namespace {

constexpr int64_t kNoAllocMaxRows = DQ4ML_NOALLOC_MAX_ROWS;
constexpr int64_t kNoAllocMaxTiles = (kNoAllocMaxRows + kFilterStateTile - 1) / kFilterStateTile;
constexpr int kNoAllocHeaderWords = 32;  // words [0,1] = the two row counters; the rest of the first 256 bytes is unused
constexpr int kNoAllocMaxDevices = 64;

// In host code this name is only a handle for the runtime, not an address a kernel could use; the device address is
// looked up once per device with cudaGetSymbolAddress (below) and that pointer is what the kernel receives.
static __device__ __align__(256) unsigned long long g_noalloc_scratch[kNoAllocHeaderWords + kNoAllocMaxTiles];

std::mutex g_noalloc_mu;
unsigned long long* g_noalloc_dev_ptr[kNoAllocMaxDevices] = {nullptr};

// Set-up step: resolves the array's address on the CURRENT device, once per device.
cudaError_t noalloc_scratch_on_current_device(unsigned long long** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kNoAllocMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_dev_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_noalloc_dev_ptr[dev] = static_cast<unsigned long long*>(p);
  }
  *out = g_noalloc_dev_ptr[dev];
  return cudaSuccess;
}

}  // namespace
}  // namespace dq4ml

extern "C" int dq4ml_dq_filter_noalloc(void* stream, const int32_t* guest, const double* price, int64_t n,
                                       const dq4ml_dq_params* params, int32_t* guest_out, double* price_out,
                                       int64_t* n_out, int64_t* n_after_rule1) {
  using namespace dq4ml;
  if (n < 0 || n_out == nullptr || (n > 0 && (guest == nullptr || price == nullptr || guest_out == nullptr || price_out == nullptr))) {
    set_error("dq_filter_noalloc: null column / output or negative n");
    return DQ4ML_ERR_INVALID;
  }
  if (n > kNoAllocMaxRows) {
    set_error("dq_filter_noalloc: %lld rows exceed the %lld-row limit of the built-in scratch (DQ4ML_NOALLOC_MAX_ROWS); "
              "split the input or use dq4ml_dq_filter",
              static_cast<long long>(n), static_cast<long long>(kNoAllocMaxRows));
    return DQ4ML_ERR_UNSUPPORTED;
  }
  if (n > 0 && (static_cast<const void*>(guest) == static_cast<const void*>(guest_out) ||
                static_cast<const void*>(price) == static_cast<const void*>(price_out))) {
    set_error("dq_filter_noalloc: in-place compaction is not supported");
    return DQ4ML_ERR_INVALID;
  }
  if ((reinterpret_cast<uintptr_t>(guest) & 3u) != 0 || (reinterpret_cast<uintptr_t>(price) & 7u) != 0 ||
      (reinterpret_cast<uintptr_t>(guest_out) & 3u) != 0 || (reinterpret_cast<uintptr_t>(price_out) & 7u) != 0) {
    set_error("dq_filter_noalloc: columns must be naturally aligned (int32: 4 bytes, double: 8 bytes)");
    return DQ4ML_ERR_INVALID;
  }
  *n_out = 0;
  if (n_after_rule1 != nullptr) *n_after_rule1 = 0;
  if (n == 0) return DQ4ML_OK;
  unsigned long long* scratch = nullptr;
  cudaError_t e = noalloc_scratch_on_current_device(&scratch);
  if (e != cudaSuccess) {
    set_error("dq_filter_noalloc: cudaGetSymbolAddress(scratch): %s", cudaGetErrorString(e));
    return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? DQ4ML_ERR_NO_DEVICE : DQ4ML_ERR_CUDA;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const dq4ml_dq_params p = params != nullptr ? *params : dq4ml_dq_params_default();
  int64_t launched = 0;
  // The array still holds the previous call's words: launch_dq_filter zeroes the counters and the first
  // ceil(n / 2048) look-back words (one per tile of kFilterStateTile rows) on `s` (cudaMemsetAsync, no allocation) before
  // the kernel reads them.
  e = launch_dq_filter(guest, price, n, p, guest_out, price_out, scratch + kNoAllocHeaderWords, scratch, s, &launched);
  unsigned long long counts[2] = {0, 0};
  if (e == cudaSuccess) e = cudaMemcpyAsync(counts, scratch, sizeof(counts), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) {
    set_error("dq_filter_noalloc: %s", cudaGetErrorString(e));
    return DQ4ML_ERR_CUDA;
  }
  *n_out = static_cast<int64_t>(counts[0]);
  if (n_after_rule1 != nullptr) *n_after_rule1 = static_cast<int64_t>(counts[1]);
  return DQ4ML_OK;
}
