// lane_group_kernels.cu -- the streaming passes of the l-bfgs / huber branches and of the evaluation over 9..12 table
// columns (weighted: 8..16 features): LeastSquaresAggregator, MultivariateOnlineSummarizer, HuberAggregator (each plain and with instance weights),
// LinearRegressionModel.transform + the summary sums, and the per-group gradient of wide tables (reached from lr.fit /
// model.evaluate, DataQuality4MachineLearningApp.java:126-139).
//
// The generic kernel (stream_reduce.cuh) keeps a whole row group in the registers of one thread.  That is the right
// shape up to 8 columns (4 rows x 8 columns x 8 B in flight per thread, 1.05 of the measured HBM peak); its 16-column
// bucket runs one CTA of ~220 registers per SM (0.75 at 9 columns, 0.88 at 12, 0.98 at 16), and a weighted fit over
// 8..15 features lands in the bucket above its feature count because the weight is one more column (0.74 at 8 weighted
// features, 0.66 at 16).  How much a pass streams is set by the bytes its SM keeps in flight (~130 KB are needed),
// i.e. by how many loaded-but-unused values the register file holds beside the accumulators, so here the columns of a
// row group are dealt out to a group of G = 1 or 2 neighbouring lanes: lane h holds columns hs h .. hs h + hs - 1
// (hs = ceil(d / G) <= 8) of PR consecutive rows.  Ops that need the row's dot product exchange the partial sums with
// a shuffle; every lane keeps the accumulators of its own columns, the scalar accumulators are kept by both lanes of a
// group and read from lane 0.  Label, weight and DQ columns are requested by both lanes at the same addresses, which
// the load unit merges.  Two register shapes: 8 columns x 4 rows at 512 threads and, for hs <= 6, 6 columns x 8 rows at
// 384 threads.  The weight of the weighted ops is a row scalar here, not column 0 of the register tile.
// One persistent CTA per SM; shuffle over the lanes of one class -> shared memory -> one partial per CTA -> the CTA
// that takes the last ticket folds the partials in a fixed order into the public layout of the generic kernel's
// bucket and publishes it (epilogue.cuh), like stream_reduce_kernel.
//
// Measured (2 x 10^8 rows, fraction of the measured HBM peak, summarizer / loss+gradient pass inside a fit,
// profiles/r2_bucket_sweep.jsonl): 9 columns 0.89 / 0.91 (generic: 0.77 / 0.75), 12: 1.00 / 0.96 (0.90 / 0.88),
// 16: 0.99 / 0.98-1.04 (1.00 / 0.98); weighted, 8 features: 1.05 / 1.07 (0.67 / 0.74), 12: 0.67 / 0.90 (0.83 / 0.88),
// 16: 0.70 / 0.97 (0.53 / 0.66).  The kernels are written for groups of 4 lanes as well (17..32 columns), but there
// round 1's lane-pair kernels (pair32_kernels.cu: 12 register columns at 384 threads up to 24 columns, groups of 4 x 8
// columns at 512 threads above) stay ahead by 3-10 % and keep those passes.  Also tried and dropped for 17..32
// columns: landing the next tiles in shared memory with cp.async so that the bytes in flight are bounded by the 227 KB
// of shared memory instead of the registers.  (a) per-thread slots, 16-byte pieces, two per 4-row column chunk, no
// barrier: a flat 0.53 of the peak whatever the width (profiles/r2_lg_ring_ldc_full.json: every 32-byte sector
// requested twice, L2 hit rate 56 %); (b) warp-private rings of 2-3 stages filled with coalesced copies (one
// instruction = two whole 256-byte column chunks), __syncwarp, conflict-free 128-bit reads, refill before the
// arithmetic, ~200 KB in flight: 0.45-0.65 (profiles/r2_lane_group_shapes.jsonl, "wc" lines).  The tcgen05 Gram
// kernel's cp.async staging stops at the same 0.65 with its loads alone: on this part LDGSTS does not stream at the
// HBM rate, whatever is in flight; the filter's cp.async.bulk ring does (0.80 with its compaction, 0.91 loads alone).
#include <cstdlib>

#include "lane_group.hpp"

namespace dq4ml {
namespace {

constexpr int kLgLsq = 0, kLgSumm = 1, kLgHuber = 2, kLgEval = 3, kLgGrad = 4;

template <int OP, bool W>
struct LgShape {
  // scalar accumulators (identical on every lane of a group) and accumulators per column
  static constexpr int NS = OP == kLgLsq ? 2 : (OP == kLgSumm ? (W ? 5 : 3) : (OP == kLgHuber ? 4 : (OP == kLgEval ? 9 : 0)));
  static constexpr int NC = OP == kLgSumm ? 2 : (OP == kLgEval ? 0 : 1);
  static constexpr bool DOT = OP == kLgLsq || OP == kLgHuber || OP == kLgEval;
};

struct LgConsts {
  double eff[kMaxCols];  // per FEATURE column: scaled coefficients (lsq, huber) or coefficients (eval)
  double s0, s1, s2;     // lsq: offset, 1 / sigma_y; huber: intercept, sigma, epsilon; eval: intercept
  int mode;              // eval: per-row output 0 = prediction, 1 = residual
  int bucket;            // column-group width of the public result layout (stream_bucket of the table)
};

template <int OP>
__host__ __device__ constexpr int lg_scalar_kind(int k) {
  return OP == kLgEval ? (k == 7 ? kMin : (k == 8 ? kMaxNan : kSum)) : kSum;
}

// public index of scalar accumulator k / of accumulator q of table column t (t counts the weight column of the
// weighted ops), in the layout of the generic Op of ops.cuh for the bucket
template <int OP, bool W>
__device__ __forceinline__ int lg_scalar_index(int k, int bucket) {
  if (OP == kLgSumm) {
    if (!W) return k == 0 ? 0 : 2 * bucket + k;           // count | label S1, S2 at 1 + 2 B, 2 + 2 B
    return k < 3 ? k : 2 * bucket + (k - 2);              // count, W, sum w^2 | label S1, S2
  }
  return k;
}
template <int OP>
__device__ __forceinline__ int lg_column_index(int q, int t) {
  if (OP == kLgLsq) return 2 + t;
  if (OP == kLgSumm) return 1 + 2 * t + q;
  if (OP == kLgHuber) return 4 + t;
  return t;  // kLgGrad
}

template <int G>
__device__ __forceinline__ double group_sum(double v) {
  if (G >= 2) v += __shfl_xor_sync(0xffffffffu, v, 1);
  if (G >= 4) v += __shfl_xor_sync(0xffffffffu, v, 2);  // (p0 + p1) + (p2 + p3) on every lane of the group
  return v;
}

// One tile of one lane in registers: PH feature columns x PR rows, the row scalars, the separate DQ columns.
template <int PH, int PR>
struct LgTile {
  double x[PH][PR];
  int g0[PR];  // raw values of feature column 0 (the lab layout keeps guest there)
  double y[PR], w[PR];
  int gs[PR];  // DQ columns when they are not feature 0 / the label
  double ps[PR];
};

// label-independent row scalars of a tile straight into registers: weight, separate DQ columns
template <bool W, int PH, int PR, bool VEC>
__device__ __forceinline__ void lg_fetch_side(LgTile<PH, PR>& t, const double* __restrict__ weight, const DqArgs& dq,
                                              const int dq_mode, const int64_t row, const int64_t n) {
  if (W) {
    if (VEC) {
      load_f64<PR, true>(weight, row, t.w);
    } else {
#pragma unroll
      for (int r = 0; r < PR; ++r) t.w[r] = row + r < n ? weight[row + r] : 0.0;
    }
  }
  if (dq_mode == kDqSeparate) {
    if (VEC) {
      load_i32<PR, true>(dq.guest, row, t.gs);
      load_f64<PR, true>(dq.price, row, t.ps);
    } else {
#pragma unroll
      for (int r = 0; r < PR; ++r) {
        t.gs[r] = row + r < n ? dq.guest[row + r] : 0;
        t.ps[r] = row + r < n ? dq.price[row + r] : 0.0;
      }
    }
  }
}

// the feature columns and the label of a tile straight into registers
template <int PH, int PR, bool VEC>
__device__ __forceinline__ void lg_fetch_direct(LgTile<PH, PR>& t, const ColTable& cols, const int d, const int hs, const int h,
                                                const double* __restrict__ label, const int64_t row, const int64_t n) {
#pragma unroll
  for (int r = 0; r < PR; ++r) {
    t.g0[r] = 0;
    t.y[r] = 0.0;
  }
#pragma unroll
  for (int j = 0; j < PH; ++j) {
    const int col = hs * h + j;
    if (j < hs && col < d) {
      if ((cols.i32_mask >> col) & 1u) {
        int v[PR];
        if (VEC) {
          load_i32<PR, true>(static_cast<const int*>(cols.ptr[col]), row, v);
        } else {
#pragma unroll
          for (int r = 0; r < PR; ++r) v[r] = row + r < n ? static_cast<const int*>(cols.ptr[col])[row + r] : 0;
        }
#pragma unroll
        for (int r = 0; r < PR; ++r) t.x[j][r] = static_cast<double>(v[r]);
        if (j == 0) {  // feature 0 on lane 0 of the group (the other lanes' copy is never used)
#pragma unroll
          for (int r = 0; r < PR; ++r) t.g0[r] = v[r];
        }
      } else if (VEC) {
        load_f64<PR, true>(static_cast<const double*>(cols.ptr[col]), row, t.x[j]);
      } else {
#pragma unroll
        for (int r = 0; r < PR; ++r) t.x[j][r] = row + r < n ? static_cast<const double*>(cols.ptr[col])[row + r] : 0.0;
      }
    } else {
#pragma unroll
      for (int r = 0; r < PR; ++r) t.x[j][r] = 0.0;
    }
  }
  if (label != nullptr) {  // the later column groups of a wide table's summarizer come without the label
    if (VEC) {
      load_f64<PR, true>(label, row, t.y);
    } else {
#pragma unroll
      for (int r = 0; r < PR; ++r) t.y[r] = row + r < n ? label[row + r] : 0.0;
    }
  }
}

// DQ predicate + the op over one tile.  s = scalar accumulators, ca = per-column accumulators.
template <int OP, bool W, int G, int PH, int PR, bool VEC>
__device__ __forceinline__ void lg_consume(double (&s)[LgShape<OP, W>::NS > 0 ? LgShape<OP, W>::NS : 1],
                                           double (&ca)[LgShape<OP, W>::NC > 0 ? LgShape<OP, W>::NC : 1][PH], const double (&cst)[PH],
                                           const double lab_c, const LgConsts& c, LgTile<PH, PR>& t, const DqArgs& dq,
                                           const int dq_mode, const int64_t row, const int64_t n, const int h,
                                           double* __restrict__ row_out) {
  using Shape = LgShape<OP, W>;
  double (&x)[PH][PR] = t.x;
  const double (&y)[PR] = t.y;
  const double (&w)[PR] = t.w;
  bool keep[PR];
  if (dq_mode == kDqSeparate) {
#pragma unroll
    for (int r = 0; r < PR; ++r) keep[r] = dq_keep(t.gs[r], t.ps[r], dq.t);
  } else if (dq_mode == kDqLab) {  // guest is feature column 0 (held by lane 0 of the group), price is the label
#pragma unroll
    for (int r = 0; r < PR; ++r) {
      const int g = G > 1 ? __shfl_sync(0xffffffffu, t.g0[r], (threadIdx.x & 31) & ~(G - 1)) : t.g0[r];
      keep[r] = dq_keep(g, y[r], dq.t);
    }
  } else {
#pragma unroll
    for (int r = 0; r < PR; ++r) keep[r] = true;
  }
  if (!VEC) {
#pragma unroll
    for (int r = 0; r < PR; ++r) keep[r] = keep[r] && row + r < n;
  }
  double out[PR];
#pragma unroll
  for (int r = 0; r < PR; ++r) {
    const bool kp = keep[r];
    const double yr = kp ? y[r] : 0.0;
    const double wr = W ? (kp ? w[r] : 0.0) : 1.0;  // a dropped row has weight 0 in every weighted sum
    if (OP != kLgGrad) {
#pragma unroll
      for (int j = 0; j < PH; ++j) x[j][r] = kp ? x[j][r] : 0.0;  // dropped rows may hold NaN / Inf payload
    }
    double total = 0.0;
    if (Shape::DOT) {
      double part = 0.0;
#pragma unroll
      for (int j = 0; j < PH; ++j) part = fma(x[j][r], cst[j], part);
      total = group_sum<G>(part);
    }
    if constexpr (OP == kLgLsq) {
      // LeastSquaresAggregator.add: diff = x.(c / sigma) - y / sigma_y + offset; lossSum += w diff^2 / 2; weightSum += w;
      // gradSum_j += w diff x_j (ops.cuh LsqOp / LsqWOp)
      const double diff = kp ? total - yr * c.s1 + c.s0 : 0.0;
      const double f = W ? wr * diff : diff;
      s[0] = fma(0.5 * f, diff, s[0]);
      s[1] += W ? wr : (kp ? 1.0 : 0.0);
#pragma unroll
      for (int j = 0; j < PH; ++j) ca[0][j] = fma(f, x[j][r], ca[0][j]);
    } else if constexpr (OP == kLgSumm) {
      // MultivariateOnlineSummarizer as shifted power sums (ops.cuh SummCoreOp / SummCoreWOp); a zero-weight row is
      // skipped entirely, a negative or NaN weight poisons the weight sum (require(weight >= 0.0))
      const bool on = W ? (kp && wr != 0.0) : kp;
      s[0] += on ? 1.0 : 0.0;
      if (W) {
        s[1] += wr >= 0.0 ? wr : __longlong_as_double(0x7ff8000000000000LL);
        s[2] = fma(wr, wr, s[2]);
      }
#pragma unroll
      for (int j = 0; j < PH; ++j) {
        const double t = on ? x[j][r] - cst[j] : 0.0;
        const double wt = W ? wr * t : t;
        ca[0][j] += wt;
        ca[1][j] = fma(wt, t, ca[1][j]);
      }
      const double t = on ? yr - lab_c : 0.0;
      const double wt = W ? wr * t : t;
      s[W ? 3 : 1] += wt;
      s[W ? 4 : 2] = fma(wt, t, s[W ? 4 : 2]);
    } else if constexpr (OP == kLgHuber) {
      // HuberAggregator.add (ops.cuh HuberOp / HuberWOp): s0 = intercept, s1 = sigma, s2 = epsilon
      const bool on = W ? (kp && wr != 0.0) : kp;
      const double margin = c.s0 + total;
      const double res = yr - margin, rs = res / c.s1;
      const bool inner = fabs(res) <= c.s1 * c.s2;
      const double loss = inner ? 0.5 * (c.s1 + res * rs) : 0.5 * (c.s1 + 2.0 * c.s2 * fabs(res) - c.s1 * c.s2 * c.s2);
      const double dsig = inner ? 0.5 * (1.0 - rs * rs) : 0.5 * (1.0 - c.s2 * c.s2);
      const double f0 = inner ? -rs : (res >= 0.0 ? -c.s2 : c.s2);
      const double f = on ? (W ? wr * f0 : f0) : 0.0;
      s[0] += on ? (W ? wr * loss : loss) : 0.0;
      s[1] += W ? (kp ? (wr >= 0.0 ? wr : __longlong_as_double(0x7ff8000000000000LL)) : 0.0) : (kp ? 1.0 : 0.0);
      s[2] += f;
      s[3] += on ? (W ? wr * dsig : dsig) : 0.0;
#pragma unroll
      for (int j = 0; j < PH; ++j) ca[0][j] = fma(f, on ? x[j][r] : 0.0, ca[0][j]);
    } else if constexpr (OP == kLgEval) {
      // LinearRegressionModel.transform + the RegressionMetrics / LinearRegressionSummary sums (ops.cuh EvalOp)
      double yhat = total + c.s0;
      yhat = kp ? yhat : 0.0;
      const double res = yr - yhat;
      s[0] += kp ? 1.0 : 0.0;
      s[1] = fma(res, res, s[1]);
      s[2] += yr;
      s[3] = fma(yr, yr, s[3]);
      s[4] += fabs(res);
      s[5] += res;
      s[6] = fma(res, yr, s[6]);
      s[7] = combine(kMin, s[7], kp ? res : __builtin_huge_val());
      s[8] = combine(kMaxNan, s[8], kp ? res : -__builtin_huge_val());
      out[r] = c.mode == 0 ? yhat : res;
    } else {
      // the second half of the LeastSquaresAggregator for wide tables (ops.cuh GradOp): gradSum_j += r x_j, where the
      // scaled difference r arrives as the label; r == 0 marks a row the rules dropped (its features may be garbage)
      const bool live = yr != 0.0;
#pragma unroll
      for (int j = 0; j < PH; ++j) ca[0][j] = fma(yr, live ? x[j][r] : 0.0, ca[0][j]);
    }
  }
  if (OP == kLgEval && row_out != nullptr && h == 0) {
    if constexpr (VEC) {
#pragma unroll
      for (int r = 0; r < PR; r += 2) stg_stream_d2(row_out + row + r, make_double2(out[r], out[r + 1]));
    } else {
#pragma unroll
      for (int r = 0; r < PR; ++r)
        if (row + r < n) row_out[row + r] = out[r];
    }
  }
}

template <int OP, bool W, int G, int PH, int PR, int BLK>
__global__ void __launch_bounds__(BLK, 1)
    lane_group_kernel(const LgConsts c, const ColTable cols, const int d, const int hs, const double* __restrict__ label,
                      const double* __restrict__ weight, const DqArgs dq, const int dq_mode, const int64_t n,
                      const double* __restrict__ dev_consts, double* __restrict__ row_out, double* __restrict__ partials,
                      unsigned int* __restrict__ ticket, double* __restrict__ result, const int nacc, const XrEpilogue ep) {
  using Shape = LgShape<OP, W>;
  constexpr int NS = Shape::NS, NC = Shape::NC, NT = NS + NC * PH;
  constexpr int kTile = (BLK / G) * PR;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, h = lane & (G - 1);
  double s[NS > 0 ? NS : 1];
  double ca[NC > 0 ? NC : 1][PH];
#pragma unroll
  for (int k = 0; k < (NS > 0 ? NS : 1); ++k) s[k] = reduce_identity(lg_scalar_kind<OP>(k));
#pragma unroll
  for (int q = 0; q < (NC > 0 ? NC : 1); ++q)
#pragma unroll
    for (int j = 0; j < PH; ++j) ca[q][j] = 0.0;
  // this lane's constants: scaled coefficients (dot-product ops) or provisional centres (summarizer; slot = table column)
  double cst[PH];
#pragma unroll
  for (int j = 0; j < PH; ++j) {
    const int col = hs * h + j;
    const bool live = j < hs && col < d;
    cst[j] = !live ? 0.0 : (OP == kLgSumm ? __ldg(dev_consts + col + (W ? 1 : 0)) : (Shape::DOT ? c.eff[col] : 0.0));
  }
  const double lab_c = OP == kLgSumm ? __ldg(dev_consts + c.bucket) : 0.0;

  const int64_t full_tiles = n / kTile;
  const int64_t in_tile = static_cast<int64_t>(threadIdx.x / G) * PR;
  for (int64_t tile = blockIdx.x; tile < full_tiles; tile += gridDim.x) {
    LgTile<PH, PR> t;
    const int64_t row = tile * kTile + in_tile;
    lg_fetch_direct<PH, PR, true>(t, cols, d, hs, h, label, row, n);
    lg_fetch_side<W, PH, PR, true>(t, weight, dq, dq_mode, row, n);
    lg_consume<OP, W, G, PH, PR, true>(s, ca, cst, lab_c, c, t, dq, dq_mode, row, n, h, row_out);
  }
  if (full_tiles * kTile < n && static_cast<int64_t>(blockIdx.x) == full_tiles % gridDim.x) {  // the ragged end, guarded loads
    LgTile<PH, PR> t;
    const int64_t row = full_tiles * kTile + in_tile;
    lg_fetch_direct<PH, PR, false>(t, cols, d, hs, h, label, row, n);
    lg_fetch_side<W, PH, PR, false>(t, weight, dq, dq_mode, row, n);
    lg_consume<OP, W, G, PH, PR, false>(s, ca, cst, lab_c, c, t, dq, dq_mode, row, n, h, row_out);
  }

  // ---- CTA reduction, the G lane classes kept apart: shuffle distances that are multiples of G stay within a class.
  // Thread accumulator e: e < NS scalars, then [q][j].
  __shared__ double s_warp[(BLK / 32) * G * NT];
  __shared__ bool s_last;
  auto acc_of = [&](int e) -> double { return e < NS ? s[e < NS ? e : 0] : ca[(e - NS) / PH][(e - NS) % PH]; };
#pragma unroll
  for (int e = 0; e < NT; ++e) {
    const int kind = e < NS ? lg_scalar_kind<OP>(e) : kSum;
    double v = acc_of(e);
#pragma unroll
    for (int off = 16; off >= G; off >>= 1) v = combine(kind, v, __shfl_down_sync(0xffffffffu, v, off));
    if (lane < G) s_warp[(warp * G + lane) * NT + e] = v;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < G * NT; e += BLK) {
    const int kind = (e % NT) < NS ? lg_scalar_kind<OP>(e % NT) : kSum;
    double v = s_warp[e];
    for (int wp = 1; wp < BLK / 32; ++wp) v = combine(kind, v, s_warp[wp * G * NT + e]);
    partials[static_cast<size_t>(blockIdx.x) * G * NT + e] = v;
  }
  // ---- grid reduction by whichever CTA finishes last, in a fixed order, into the generic kernel's public layout
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int k = threadIdx.x; k < nacc; k += BLK) result[k] = 0.0;  // columns >= d, the weight's slots
  __syncthreads();
  __shared__ double s_red[BLK / 32];
  for (int e = 0; e < G * NT; ++e) {
    const int cls = e / NT, k = e % NT;
    int out_index, kind = kSum;
    if (k < NS) {
      if (cls != 0) continue;  // every lane of a group holds the same scalars
      kind = lg_scalar_kind<OP>(k);
      out_index = lg_scalar_index<OP, W>(k, c.bucket);
    } else {
      const int q = (k - NS) / PH, j = (k - NS) % PH, col = hs * cls + j;
      if (j >= hs || col >= d) continue;
      out_index = lg_column_index<OP>(q, col + (W ? 1 : 0));
    }
    double v = reduce_identity(kind);
    for (unsigned b = threadIdx.x; b < gridDim.x; b += BLK) v = combine(kind, v, __ldcg(partials + static_cast<size_t>(b) * G * NT + e));
    v = block_reduce(v, kind, s_red);
    if (threadIdx.x == 0) result[out_index] = v;
  }
  if (threadIdx.x == 0) *ticket = 0u;  // ready for the next launch on this stream
  xr_publish(ep, result, nacc, dev_consts, [](int k) { return lg_scalar_kind<OP>(k); });  // scalar slots of eval are 0..8 in place
}

struct LgArgs {
  LgConsts c;
  ColTable cols;  // feature columns only
  int d;
  const double* label;
  const double* weight;
  DqArgs dq;
  int dq_mode;
  int64_t n;
  double* row_out;
  int nacc;
};

template <int OP, bool W, int G, int PH, int PR, int BLK>
cudaError_t launch_shape(const LgArgs& a, const Scratch& scratch, cudaStream_t stream) {
  constexpr int NT = LgShape<OP, W>::NS + LgShape<OP, W>::NC * PH;
  constexpr int kTile = (BLK / G) * PR;
  const int64_t tiles = (a.n + kTile - 1) / kTile;
  int64_t grid = scratch.sm_count;
  if (grid > tiles) grid = tiles;
  if (grid < 1) grid = 1;
  if (static_cast<size_t>(grid) * G * NT * sizeof(double) > scratch.partials_bytes) return cudaErrorMemoryAllocation;
  const int hs = (a.d + G - 1) / G;  // columns per lane: the table is dealt out evenly over the lane group
  lane_group_kernel<OP, W, G, PH, PR, BLK><<<static_cast<unsigned>(grid), BLK, 0, stream>>>(
      a.c, a.cols, a.d, hs, a.label, a.weight, a.dq, a.dq_mode, a.n, scratch.consts, a.row_out, scratch.partials, scratch.ticket,
      scratch.result, a.nacc, scratch.ep);
  return cudaGetLastError();
}

// G = the smallest group that leaves at most 8 columns per lane; register shape by the columns per lane
template <int OP, bool W, int G>
cudaError_t launch_group(const LgArgs& a, const Scratch& scratch, cudaStream_t stream) {
  const int hs = (a.d + G - 1) / G;
  // the tall shape for the unweighted ops (summarizer at 9 / 12 columns: 0.89 / 1.00 against 0.76 / 0.90 with the flat
  // one); with the weight's registers on top it spills and the flat shape is ahead (loss/gradient at 12 weighted
  // features: 0.96 against 0.92)
  if (!W && hs <= 6) return launch_shape<OP, W, G, 6, 8, 384>(a, scratch, stream);
  return launch_shape<OP, W, G, 8, 4, 512>(a, scratch, stream);
}

template <int OP, bool W>
cudaError_t launch_op(const LgArgs& a, const Scratch& scratch, cudaStream_t stream) {
  if (a.d <= 8) {
    if constexpr (W) return launch_group<OP, W, 1>(a, scratch, stream);
    return cudaErrorInvalidValue;  // up to 8 unweighted columns are the generic kernel's
  }
  if (a.d <= 16) return launch_group<OP, W, 2>(a, scratch, stream);
  if constexpr (!W && (OP == kLgHuber || OP == kLgEval || OP == kLgGrad)) return launch_group<OP, W, 4>(a, scratch, stream);
  return cudaErrorInvalidValue;  // the other ops at 17..32 columns: pair32_kernels.cu / the generic 32-column bucket
}

}  // namespace

bool lane_group_takes(StreamOpKind op, int table_cols) {
  static const char* off = std::getenv("DQ4ML_LANE_GROUPS");  // DQ4ML_LANE_GROUPS=0: the generic kernels everywhere (A/B)
  if (off != nullptr && off[0] == '0') return false;
  switch (op) {
    // measured against the generic kernel of the same bucket, op by op (profiles/r2_bucket_sweep*.jsonl):
    case kOpHuber: case kOpEval: case kOpGrad:
      if (table_cols >= 17 && table_cols <= kMaxCols) {  // groups of four lanes x 8 columns at 512 threads: evaluation 0.81 / 0.91 at 24 / 32
                                                         // columns against 0.77 / 0.81 in the generic 32-column bucket, huber level
        static const char* quad = std::getenv("DQ4ML_LG_QUAD");
        return quad == nullptr || quad[0] != '0';
      }
      return table_cols >= 9 && table_cols <= 12;
    case kOpLsq: case kOpSummCore:
      return table_cols >= 9 && table_cols <= 12;   // 0.88-0.96 against 0.76-0.90; from 13 columns the generic bucket is level
    case kOpLsqW: case kOpHuberW:
      return table_cols >= 9 && table_cols <= 17;   // the weight + 8..16 features: 0.85-1.07 against 0.66-0.95
    case kOpSummCoreW:
      return table_cols == 9 || table_cols == 17;   // where the weight pushes the generic kernel into the next bucket: 1.05 / 0.72 against 0.68 / 0.53
    default: return false;
  }
}

cudaError_t launch_lane_group(StreamOpKind op, const ColTable& cols, int d, const double* label, const DqArgs& dq, int dq_mode,
                              int64_t n, const void* host_consts, double* row_out, const Scratch& scratch, cudaStream_t stream) {
  const bool weighted = op == kOpLsqW || op == kOpSummCoreW || op == kOpHuberW;
  LgArgs a{};
  a.c.bucket = stream_bucket(d);
  a.nacc = stream_nacc(op, a.c.bucket);
  const int first = weighted ? 1 : 0;  // the weight is column 0 of a weighted op's table
  a.d = d - first;
  if (a.d < 1 || a.d > kMaxCols) return cudaErrorInvalidValue;
  for (int j = 0; j < kMaxStreamD; ++j) a.cols.ptr[j] = j < a.d ? cols.ptr[j + first] : nullptr;
  a.cols.i32_mask = cols.i32_mask >> first;
  a.weight = weighted ? static_cast<const double*>(cols.ptr[0]) : nullptr;
  if (weighted && (cols.i32_mask & 1u)) return cudaErrorInvalidValue;  // the weight column is fp64
  a.label = label;
  a.dq = dq;
  a.dq_mode = dq_mode;
  a.n = n;
  a.row_out = row_out;
  switch (op) {
    case kOpLsq: case kOpLsqW: {
      if (host_consts == nullptr) return cudaErrorInvalidValue;
      const LsqConsts* k = static_cast<const LsqConsts*>(host_consts);
      for (int j = 0; j < a.d; ++j) a.c.eff[j] = k->eff[j + first];
      a.c.s0 = k->offset;
      a.c.s1 = k->inv_ystd;
      return weighted ? launch_op<kLgLsq, true>(a, scratch, stream) : launch_op<kLgLsq, false>(a, scratch, stream);
    }
    case kOpHuber: case kOpHuberW: {
      if (host_consts == nullptr) return cudaErrorInvalidValue;
      const HuberConsts* k = static_cast<const HuberConsts*>(host_consts);
      for (int j = 0; j < a.d; ++j) a.c.eff[j] = k->eff[j + first];
      a.c.s0 = k->intercept;
      a.c.s1 = k->sigma;
      a.c.s2 = k->epsilon;
      return weighted ? launch_op<kLgHuber, true>(a, scratch, stream) : launch_op<kLgHuber, false>(a, scratch, stream);
    }
    case kOpEval: {
      if (host_consts == nullptr) return cudaErrorInvalidValue;
      const EvalConsts* k = static_cast<const EvalConsts*>(host_consts);
      for (int j = 0; j < a.d; ++j) a.c.eff[j] = k->coef[j];
      a.c.s0 = k->intercept;
      a.c.mode = k->mode;
      return launch_op<kLgEval, false>(a, scratch, stream);
    }
    case kOpSummCore: return launch_op<kLgSumm, false>(a, scratch, stream);
    case kOpSummCoreW: return launch_op<kLgSumm, true>(a, scratch, stream);
    case kOpGrad: return launch_op<kLgGrad, false>(a, scratch, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace dq4ml
