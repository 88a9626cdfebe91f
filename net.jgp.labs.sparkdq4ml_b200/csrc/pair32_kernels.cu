// pair32_kernels.cu -- the summarizer and loss/gradient passes of the l-bfgs branch for 17..32 feature columns
// (MultivariateOnlineSummarizer / LeastSquaresAggregator reached from lr.fit, App:126, solver l-bfgs).
//
// The generic streaming kernel keeps a whole row group in registers: at 32 columns that is 2 rows per thread, 128-bit
// loads and ~250 registers, and it measured 0.78 of the HBM roofline, against 0.99 for the 16-column bucket with 4 rows
// per thread and 256-bit loads.  Here the columns of a row group are split evenly over a group of G = 2 or 4
// neighbouring lanes: lane h holds columns hs h .. hs h + hs - 1, hs = ceil(d / G), of 4 consecutive rows, the partial
// dot products of a row are exchanged with one or two shuffles, and each lane keeps the gradient (or the power sums)
// of its own columns.  Label and DQ columns are requested by all lanes of a group at the same addresses, which the
// load unit merges.  One persistent CTA per SM; warp shuffle over the lanes of one class -> shared memory -> one
// partial per CTA -> the CTA that takes the last ticket folds the partials in a fixed order, like stream_reduce_kernel.
//
// A pass is bound by how many loaded-but-unused bytes the registers of an SM can hold in flight (~130-150 KB are
// needed).  Round 1 (lane pairs only; 4 x 10^8 rows, one loss/gradient pass timed alone): 12 register columns at 384
// threads up to 24 columns 0.92-1.02, 16 register columns at 256 threads (230 registers, 8 warps) 0.90 at 32 columns.
// Round 2 added the groups of four lanes (8 register columns at 512 threads for 25..32 columns: 0.93 / 1.00 / 1.04 at
// 25 / 28 / 32 columns timed alone, 0.85-0.92 inside a fit where the lane pairs gave 0.82-0.88; ncu of the lane pair
// at 32 columns: profiles/r2_pair32_lsq_d32_full.json) and picks the shape per op from profiles/r2_bucket_sweep.jsonl.
// Tried instead in round 1: a 3-stage shared-memory ring (256-row tiles, one cp.async.bulk of 2 KB per column slot on an
// mbarrier, one row per thread out of shared memory into the generic Op bodies).  Its loads alone reach 0.91, but the
// one-row-per-thread consumer is a serial fp64 chain (the 32-term dot product) on 8 warps and the whole pass measured
// 0.45 (0.52 with 16-byte cp.async pieces).
#include <cstdlib>

#include "pair32.hpp"

namespace dq4ml {
namespace {

constexpr int PR = 4;    // rows per lane pair and tile
// PH = register columns per lane, G = lanes per row group, BLK = threads per CTA; a tile is BLK / G x PR rows.

template <int OP, int PH>
struct PairShape {
  static constexpr int NT = OP == kPairLsq ? 2 + PH : 1 + 2 * PH + 2;  // accumulators per thread
};

// index of a thread accumulator in the public result vector of the 32-column bucket, or -1 if this lane class does not
// own it (LsqOp<32>: loss, count, grad[32]; SummCoreOp<32>: count, (S1, S2) x 32 columns, (S1, S2) of the label)
template <int OP, int PH>
__device__ __forceinline__ int final_index(int h, int k, int hs, int d) {
  if (OP == kPairLsq) {
    if (k < 2) return h == 0 ? k : -1;
    const int j = k - 2, c = hs * h + j;
    return (j < hs && c < d) ? 2 + c : -1;
  }
  if (k == 0) return h == 0 ? 0 : -1;
  if (k <= 2 * PH) {
    const int j = (k - 1) >> 1, c = hs * h + j;
    return (j < hs && c < d) ? 1 + 2 * c + ((k - 1) & 1) : -1;
  }
  return h == 0 ? 1 + 2 * 32 + (k - 1 - 2 * PH) : -1;
}

template <int OP, int DQMODE, bool VEC, int PH, int G>
__device__ __forceinline__ void pair_rows(double (&acc)[PairShape<OP, PH>::NT], const double (&cst)[PH], const double lab_c,
                                          const double offset, const double inv_ystd, const ColTable& cols, const int d,
                                          const int hs, const double* __restrict__ label, const DqArgs& dq,
                                          const int64_t row, const int64_t n, const int h) {
  double x[PH][PR];
  int g0[PR];  // raw values of column 0 (the lab layout keeps guest there)
  double y[PR];
#pragma unroll
  for (int r = 0; r < PR; ++r) {
    g0[r] = 0;
    y[r] = 0.0;
  }
#pragma unroll
  for (int j = 0; j < PH; ++j) {
    const int c = hs * h + j;
    if (j < hs && c < d) {
      if ((cols.i32_mask >> c) & 1u) {
        int t[PR];
        if (VEC) {
          load_i32<PR, true>(static_cast<const int*>(cols.ptr[c]), row, t);
        } else {
#pragma unroll
          for (int r = 0; r < PR; ++r) t[r] = row + r < n ? static_cast<const int*>(cols.ptr[c])[row + r] : 0;
        }
#pragma unroll
        for (int r = 0; r < PR; ++r) x[j][r] = static_cast<double>(t[r]);
        if (j == 0) {  // column 0 for the even lane (the odd lane's copy is never used)
#pragma unroll
          for (int r = 0; r < PR; ++r) g0[r] = t[r];
        }
      } else if (VEC) {
        load_f64<PR, true>(static_cast<const double*>(cols.ptr[c]), row, x[j]);
      } else {
#pragma unroll
        for (int r = 0; r < PR; ++r) x[j][r] = row + r < n ? static_cast<const double*>(cols.ptr[c])[row + r] : 0.0;
      }
    } else {
#pragma unroll
      for (int r = 0; r < PR; ++r) x[j][r] = 0.0;
    }
  }
  if (label != nullptr) {  // the later column groups of a wide table's summarizer come without the label
    if (VEC) {
      load_f64<PR, true>(label, row, y);
    } else {
#pragma unroll
      for (int r = 0; r < PR; ++r) y[r] = row + r < n ? label[row + r] : 0.0;
    }
  }
  bool keep[PR];
  if (DQMODE == kDqSeparate) {
    int g[PR];
    double pr[PR];
    if (VEC) {
      load_i32<PR, true>(dq.guest, row, g);
      load_f64<PR, true>(dq.price, row, pr);
    } else {
#pragma unroll
      for (int r = 0; r < PR; ++r) {
        g[r] = row + r < n ? dq.guest[row + r] : 0;
        pr[r] = row + r < n ? dq.price[row + r] : 0.0;
      }
    }
#pragma unroll
    for (int r = 0; r < PR; ++r) keep[r] = dq_keep(g[r], pr[r], dq.t);
  } else if (DQMODE == kDqLab) {  // guest is feature column 0 (held by the even lane), price is the label
#pragma unroll
    for (int r = 0; r < PR; ++r) {
      const int g = __shfl_sync(0xffffffffu, g0[r], (threadIdx.x & 31) & ~(G - 1));
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
#pragma unroll
  for (int r = 0; r < PR; ++r) {
    const double yr = keep[r] ? y[r] : 0.0;
    if (OP == kPairLsq) {
      double part = 0.0;
#pragma unroll
      for (int j = 0; j < PH; ++j) {
        x[j][r] = keep[r] ? x[j][r] : 0.0;  // dropped rows may hold NaN / Inf payload
        part = fma(x[j][r], cst[j], part);
      }
      double dot = part + __shfl_xor_sync(0xffffffffu, part, 1);
      if (G == 4) dot += __shfl_xor_sync(0xffffffffu, dot, 2);  // (p0 + p1) + (p2 + p3) on every lane of the group
      const double diff = keep[r] ? dot - yr * inv_ystd + offset : 0.0;
      if (h == 0) {
        acc[0] = fma(0.5 * diff, diff, acc[0]);
        acc[1] += keep[r] ? 1.0 : 0.0;
      }
#pragma unroll
      for (int j = 0; j < PH; ++j) acc[2 + j] = fma(diff, x[j][r], acc[2 + j]);
    } else {
      if (h == 0) acc[0] += keep[r] ? 1.0 : 0.0;
#pragma unroll
      for (int j = 0; j < PH; ++j) {
        const double t = keep[r] ? x[j][r] - cst[j] : 0.0;
        acc[1 + 2 * j] += t;
        acc[2 + 2 * j] = fma(t, t, acc[2 + 2 * j]);
      }
      if (h == 0) {
        const double t = keep[r] ? yr - lab_c : 0.0;
        acc[1 + 2 * PH] += t;
        acc[2 + 2 * PH] = fma(t, t, acc[2 + 2 * PH]);
      }
    }
  }
}

struct PairConsts {
  double eff[32];
  double offset, inv_ystd;
};

template <int OP, int DQMODE, int PH, int BLK, int G>
__global__ void __launch_bounds__(BLK, 1)
    pair32_kernel(const PairConsts c, const ColTable cols, const int d, const double* __restrict__ label, const DqArgs dq,
                  const int64_t n, const int hs, const double* __restrict__ centres, double* __restrict__ partials,
                  unsigned int* __restrict__ ticket, double* __restrict__ result, const XrEpilogue ep) {
  constexpr int NT = PairShape<OP, PH>::NT;
  constexpr int kPairTile = (BLK / G) * PR;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, h = lane & (G - 1);
  double acc[NT];
#pragma unroll
  for (int k = 0; k < NT; ++k) acc[k] = 0.0;
  double cst[PH];  // this lane's scaled coefficients (loss/gradient) or provisional centres (summarizer)
#pragma unroll
  for (int j = 0; j < PH; ++j) {
    const int col = hs * h + j;
    cst[j] = (j < hs && col < 32) ? (OP == kPairLsq ? c.eff[col] : __ldg(centres + col)) : 0.0;
  }
  const double lab_c = OP == kPairLsq ? 0.0 : __ldg(centres + 32);

  const int64_t full_tiles = n / kPairTile;
  const int64_t in_tile = static_cast<int64_t>(threadIdx.x / G) * PR;
  for (int64_t tile = blockIdx.x; tile < full_tiles; tile += gridDim.x)
    pair_rows<OP, DQMODE, true, PH, G>(acc, cst, lab_c, c.offset, c.inv_ystd, cols, d, hs, label, dq, tile * kPairTile + in_tile, n, h);
  if (full_tiles * kPairTile < n && static_cast<int64_t>(blockIdx.x) == full_tiles % gridDim.x)  // the ragged end, guarded loads
    pair_rows<OP, DQMODE, false, PH, G>(acc, cst, lab_c, c.offset, c.inv_ystd, cols, d, hs, label, dq, full_tiles * kPairTile + in_tile, n, h);

  // ---- CTA reduction, the G lane classes kept apart: shuffle distances that are multiples of G stay within a class
  __shared__ double s_warp[(BLK / 32) * G * NT];
  __shared__ bool s_last;
#pragma unroll
  for (int k = 0; k < NT; ++k) {
    double v = acc[k];
#pragma unroll
    for (int off = 16; off >= G; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
    if (lane < G) s_warp[(warp * G + lane) * NT + k] = v;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < G * NT; e += BLK) {
    double v = s_warp[e];
    for (int w = 1; w < BLK / 32; ++w) v += s_warp[w * G * NT + e];
    partials[static_cast<size_t>(blockIdx.x) * G * NT + e] = v;
  }
  // ---- grid reduction by whichever CTA finishes last, in a fixed order
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  __shared__ double s_red[BLK / 32];
  for (int e = 0; e < G * NT; ++e) {
    const int out = final_index<OP, PH>(e / NT, e % NT, hs, d);
    if (out < 0) continue;  // uniform over the CTA
    double v = 0.0;
    for (unsigned b = threadIdx.x; b < gridDim.x; b += BLK) v += __ldcg(partials + static_cast<size_t>(b) * G * NT + e);
    v = block_reduce(v, kSum, s_red);
    if (threadIdx.x == 0) result[out] = v;
  }
  if (threadIdx.x == 0) *ticket = 0u;  // ready for the next launch on this stream
  // the vector has the 32-column bucket's public layout (LsqOp<32>: 34 sums; SummCoreOp<32>: 67 sums + 33 centres)
  xr_publish(ep, result, OP == kPairLsq ? 2 + 32 : 1 + 2 * 33, centres, [](int) { return static_cast<int>(kSum); });
}

template <int OP, int PH, int BLK, int G>
cudaError_t launch_shape(int dq_mode, const PairConsts& c, const ColTable& cols, int d, const double* label, const DqArgs& dq,
                         int64_t n, const Scratch& scratch, cudaStream_t stream) {
  constexpr int kPairTile = (BLK / G) * PR;
  const int64_t tiles = (n + kPairTile - 1) / kPairTile;
  int64_t grid = scratch.sm_count;
  if (grid > tiles) grid = tiles;
  if (grid < 1) grid = 1;
  if (static_cast<size_t>(grid) * G * PairShape<OP, PH>::NT * sizeof(double) > scratch.partials_bytes) return cudaErrorMemoryAllocation;
  const unsigned gsz = static_cast<unsigned>(grid);
  const int hs = (d + G - 1) / G;  // columns per lane: the table is split evenly over the lane group
  if (dq_mode == kDqNone)
    pair32_kernel<OP, kDqNone, PH, BLK, G><<<gsz, BLK, 0, stream>>>(c, cols, d, label, dq, n, hs, scratch.consts, scratch.partials, scratch.ticket, scratch.result, scratch.ep);
  else if (dq_mode == kDqSeparate)
    pair32_kernel<OP, kDqSeparate, PH, BLK, G><<<gsz, BLK, 0, stream>>>(c, cols, d, label, dq, n, hs, scratch.consts, scratch.partials, scratch.ticket, scratch.result, scratch.ep);
  else
    pair32_kernel<OP, kDqLab, PH, BLK, G><<<gsz, BLK, 0, stream>>>(c, cols, d, label, dq, n, hs, scratch.consts, scratch.partials, scratch.ticket, scratch.result, scratch.ep);
  return cudaGetLastError();
}

template <int OP>
cudaError_t launch_mode(int dq_mode, const PairConsts& c, const ColTable& cols, int d, const double* label, const DqArgs& dq,
                        int64_t n, const Scratch& scratch, cudaStream_t stream) {
  // Shapes by measurement (profiles/r2_bucket_sweep.jsonl, r2_lane_group_shapes.jsonl; fractions of the measured HBM peak
  // inside a fit).  Up to 24 columns: the loss/gradient pass streams best as lane pairs of 12 register columns at 384
  // threads (0.89-0.96; groups of four lanes x 6 columns at 512 threads: 0.74-0.82), the summarizer the other way round
  // (0.82-0.96 against 0.79-0.82).  25..32 columns: groups of four lanes x 8 columns at 512 threads (0.85-0.92 and
  // 0.88-0.98; lane pairs of 16 columns at 256 threads, 230 registers: 0.82-0.88; at 320 threads ptxas spills 350..630
  // bytes per thread: 0.73).
  if (d <= 24) {
    if (OP == kPairSumm) return launch_shape<OP, 6, 512, 4>(dq_mode, c, cols, d, label, dq, n, scratch, stream);
    return launch_shape<OP, 12, 384, 2>(dq_mode, c, cols, d, label, dq, n, scratch, stream);
  }
  return launch_shape<OP, 8, 512, 4>(dq_mode, c, cols, d, label, dq, n, scratch, stream);
}

}  // namespace

cudaError_t launch_pair32(int op, const ColTable& cols, int d, const double* label, const DqArgs& dq, int dq_mode, int64_t n,
                          const LsqConsts* lsq, const Scratch& scratch, cudaStream_t stream) {
  PairConsts c{};
  if (op == kPairLsq) {
    if (lsq == nullptr) return cudaErrorInvalidValue;
    for (int j = 0; j < 32; ++j) c.eff[j] = lsq->eff[j];
    c.offset = lsq->offset;
    c.inv_ystd = lsq->inv_ystd;
    return launch_mode<kPairLsq>(dq_mode, c, cols, d, label, dq, n, scratch, stream);
  }
  return launch_mode<kPairSumm>(dq_mode, c, cols, d, label, dq, n, scratch, stream);
}

}  // namespace dq4ml
