// stream_reduce.cuh -- the one streaming pass every aggregation of the path is an instance of.
//
// A pass reads each column of the VectorAssembler table (plus label, plus the DQ columns when they are not already
// among them) exactly once with 128-bit loads, applies the DQ predicate in registers, feeds each surviving row to an
// `Op` that keeps NACC fp64 accumulators per thread, and ends with warp-shuffle -> shared-memory -> one partial per
// CTA -> last-CTA fixed-order reduction.  No atomics on data, so results are reproducible for a given (n, grid).
//
// This is the GPU counterpart of `instances.treeAggregate(zero)(seqOp, combOp)` in Spark ML 2.4.4, which is what
// lr.fit(df) (DataQuality4MachineLearningApp.java:126) spends its time in.
#pragma once

#include "device_utils.cuh"
#include "epilogue.cuh"

namespace dq4ml {

struct ColTable {
  const void* ptr[kMaxStreamD];
  uint32_t i32_mask;  // bit j: column j holds int32 (else fp64)
};

struct DqArgs {
  const int* guest;
  const double* price;
  DqThresholds t;
};

// DQ fusion modes
constexpr int kDqNone = 0;      // every row counts
constexpr int kDqSeparate = 1;  // guest/price are loaded as extra columns
constexpr int kDqLab = 2;       // the lab's layout: guest is feature column 0 (int32) and price is the label -> no extra bytes

// DENSE: what the table looks like is known at compile time -- 0 = nothing (d and the column types are tested per
// column and tile); otherwise all DMAX columns are there and DENSE - 1 is the bit mask of the int32 columns: 1 = all
// fp64, 2 = column 0 is int32 (the lab's layout: guest first), 3 = column 1 is int32 (weighted: weight, guest, ...).
// The tests of the general form put ~14 instructions between two loads of a tile; with them gone a thread issues its
// 17 loads back to back: 10^9 rows x 16 columns 21.0 -> 18.3 ms per pass, 0.96 -> 1.10 of the measured copy peak
// (tools/k4_variants.cu reaches the same 1.11 with this shape outside the library, a pure read 1.10-1.11).
// 5, 6, 7: the same column types (mask DENSE - 5) with fewer than DMAX columns: the width test stays, the type tests go
// (loss/gradient at 12 / 14 / 28 columns: 1.05 / 1.08 / 1.06 against 0.94-1.01 / 0.93 / 0.89-0.96 on the lane-group
// and lane-pair kernels; the width test costs the summarizer more than it gains, which keeps those kernels).
template <class Op, int DMAX, int ROWS, int DQMODE, bool VEC, int DENSE = 0>
__device__ __forceinline__ void process_rows(const Op& op, double (&acc)[Op::NACC], const double* consts,
                                             const ColTable& cols, int d,
                                             const double* __restrict__ label, const DqArgs& dq, int64_t row,
                                             double* __restrict__ row_out) {
  double x[DMAX][ROWS];
  int g[ROWS];
  double y[ROWS];
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
    g[r] = 0;
    y[r] = 0.0;
  }
#pragma unroll
  for (int j = 0; j < DMAX; ++j) {
    if ((DENSE >= 1 && DENSE <= 3) || j < d) {
      if (DENSE == 0 ? ((cols.i32_mask >> j) & 1u) != 0u : (((((DENSE - 1) & 3) >> j) & 1) != 0)) {
        int t[ROWS];
        load_i32<ROWS, VEC>(static_cast<const int*>(cols.ptr[j]), row, t);
#pragma unroll
        for (int r = 0; r < ROWS; ++r) x[j][r] = static_cast<double>(t[r]);
        if (DQMODE == kDqLab && j == 0) {
#pragma unroll
          for (int r = 0; r < ROWS; ++r) g[r] = t[r];
        }
      } else {
        load_f64<ROWS, VEC>(static_cast<const double*>(cols.ptr[j]), row, x[j]);
      }
    } else {
#pragma unroll
      for (int r = 0; r < ROWS; ++r) x[j][r] = 0.0;
    }
  }
  if (label != nullptr) load_f64<ROWS, VEC>(label, row, y);
  bool keep[ROWS];
  if (DQMODE == kDqSeparate) {
    double pr[ROWS];
    load_i32<ROWS, VEC>(dq.guest, row, g);
    load_f64<ROWS, VEC>(dq.price, row, pr);
#pragma unroll
    for (int r = 0; r < ROWS; ++r) keep[r] = dq_keep(g[r], pr[r], dq.t);
  } else if (DQMODE == kDqLab) {
#pragma unroll
    for (int r = 0; r < ROWS; ++r) keep[r] = dq_keep(g[r], y[r], dq.t);
  } else {
#pragma unroll
    for (int r = 0; r < ROWS; ++r) keep[r] = true;
  }
  double o[ROWS];
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
    double xr[DMAX];  // dropped rows reach the Op as all-zero rows (their payload may be NaN/Inf garbage)
#pragma unroll
    for (int j = 0; j < DMAX; ++j) xr[j] = keep[r] ? x[j][r] : 0.0;
    o[r] = op.row(acc, consts, xr, keep[r] ? y[r] : 0.0, keep[r], d);
  }
  if (Op::HAS_OUT && row_out != nullptr) {
    if constexpr (VEC && ROWS >= 2) {
#pragma unroll
      for (int r = 0; r < ROWS; r += 2) stg_stream_d2(row_out + row + r, make_double2(o[r], o[r + 1]));
    } else {
#pragma unroll
      for (int r = 0; r < ROWS; ++r) row_out[row + r] = o[r];
    }
  }
}

// CTA reduction (shuffle inside each warp, one shared-memory hop across the 8 warps), one partial per CTA, grid
// reduction in a fixed order by whichever CTA finishes last, hand-over of the result (epilogue.cuh).
template <class Op>
__device__ __forceinline__ void reduce_and_publish(double (&acc)[Op::NACC], const double* __restrict__ dev_consts,
                                                   double* __restrict__ partials, unsigned int* __restrict__ ticket,
                                                   double* __restrict__ result, const XrEpilogue& ep) {
  constexpr int NACC = Op::NACC;
  __shared__ double s_warp[(kBlock / 32) * NACC];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NACC; ++k) {
    const double v = warp_reduce(acc[k], Op::kind(k));
    if (lane == 0) s_warp[warp * NACC + k] = v;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < NACC; k += kBlock) {
    double v = s_warp[k];
    for (int w = 1; w < kBlock / 32; ++w) v = combine(Op::kind(k), v, s_warp[w * NACC + k]);
    partials[static_cast<size_t>(blockIdx.x) * NACC + k] = v;
  }
  // ---- grid reduction by whichever CTA finishes last, in a fixed order
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  __shared__ double s_red[kBlock / 32];
  for (int k = 0; k < NACC; ++k) {
    const int kind = Op::kind(k);
    double v = Op::identity(k);
    for (unsigned b = threadIdx.x; b < gridDim.x; b += kBlock)
      v = combine(kind, v, __ldcg(partials + static_cast<size_t>(b) * NACC + k));
    v = block_reduce(v, kind, s_red);
    if (threadIdx.x == 0) result[k] = v;
  }
  if (threadIdx.x == 0) *ticket = 0u;  // ready for the next launch on this stream
  xr_publish(ep, result, NACC, dev_consts, [](int k) { return Op::kind(k); });
}

template <class Op, int DMAX, int ROWS, int DQMODE, int DENSE = 0>
__global__ void __launch_bounds__(kBlock, (DMAX * ROWS >= 64 ? 1 : (DMAX <= 2 ? 4 : (DMAX == 4 ? 3 : 2))))
    stream_reduce_kernel(const Op op, const ColTable cols, const int d, const double* __restrict__ label,
                         const DqArgs dq, const int64_t n, const double* __restrict__ dev_consts,
                         double* __restrict__ row_out, double* __restrict__ partials, unsigned int* __restrict__ ticket,
                         double* __restrict__ result, const XrEpilogue ep) {
  constexpr int NACC = Op::NACC;
  double acc[NACC];
#pragma unroll
  for (int k = 0; k < NACC; ++k) acc[k] = Op::identity(k);
  double consts[Op::NCONST > 0 ? Op::NCONST : 1];
#pragma unroll
  for (int k = 0; k < Op::NCONST; ++k) consts[k] = __ldg(dev_consts + k);

  // Full tiles first, with no tail test inside the loop so that it can be unrolled and the loads of the next tile are
  // in flight during the arithmetic of this one (the ragged end of the table is one extra partial tile below).
  const int64_t tile_rows = static_cast<int64_t>(kBlock) * ROWS;
  const int64_t full_tiles = n / tile_rows;
#pragma unroll (DMAX >= 32 ? 1 : 2)
  for (int64_t tile = blockIdx.x; tile < full_tiles; tile += gridDim.x) {
    const int64_t row = tile * tile_rows + static_cast<int64_t>(threadIdx.x) * ROWS;
    process_rows<Op, DMAX, ROWS, DQMODE, true, DENSE>(op, acc, consts, cols, d, label, dq, row, row_out);
  }
  if (full_tiles * tile_rows < n && static_cast<int64_t>(blockIdx.x) == full_tiles % gridDim.x) {
    const int64_t row = full_tiles * tile_rows + static_cast<int64_t>(threadIdx.x) * ROWS;
    if (row + ROWS <= n) {
      process_rows<Op, DMAX, ROWS, DQMODE, true, DENSE>(op, acc, consts, cols, d, label, dq, row, row_out);
    } else {
      for (int64_t r = row; r < n; ++r) process_rows<Op, DMAX, 1, DQMODE, false>(op, acc, consts, cols, d, label, dq, r, row_out);
    }
  }

  reduce_and_publish<Op>(acc, dev_consts, partials, ticket, result, ep);
}

// The same pass for the lab's own shape -- ONE int32 feature column (guest) + the label, DQ in the lab layout or off --
// software-pipelined: every thread holds NV 8-row vectors of the tile it works on and has the NV vectors of its next
// tile in flight while it does the arithmetic, one 256-thread CTA per SM.  tools/k2_variants.cu (10^8 rows of 12 B):
// the grid-strided shape above 0.187-0.190 ms at its best occupancy, this one 0.174-0.175 ms with 24 rows per thread
// (a pure read of the same bytes: 0.170-0.172 ms).  Fewer, fatter threads: the bytes in flight (2 x 24 rows x 12 B x
// 256 threads = 147 KB per SM) do not depend on the warp scheduler finding another warp to run.
template <class Op, int NV, int DQMODE>
__global__ void __launch_bounds__(kBlock, 1)
    stream_pipe_kernel(const Op op, const int* __restrict__ col0, const double* __restrict__ label, const DqThresholds t,
                       const int64_t n, const double* __restrict__ dev_consts, double* __restrict__ row_out,
                       double* __restrict__ partials, unsigned int* __restrict__ ticket, double* __restrict__ result,
                       const XrEpilogue ep) {
  static_assert(DQMODE == kDqNone || DQMODE == kDqLab, "separate DQ columns take the grid-strided kernel");
  constexpr int NACC = Op::NACC;
  double acc[NACC];
#pragma unroll
  for (int k = 0; k < NACC; ++k) acc[k] = Op::identity(k);
  double consts[Op::NCONST > 0 ? Op::NCONST : 1];
#pragma unroll
  for (int k = 0; k < Op::NCONST; ++k) consts[k] = __ldg(dev_consts + k);

  constexpr int64_t kTile = static_cast<int64_t>(kBlock) * 8 * NV;
  const int64_t full_tiles = n / kTile;
  auto load = [&](int64_t tile, int (&g)[NV][8], double (&y)[NV][8]) {
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const int64_t row = tile * kTile + (static_cast<int64_t>(v) * kBlock + threadIdx.x) * 8;
      load_i32<8, true>(col0, row, g[v]);
      if (label != nullptr) {
        load_f64<8, true>(label, row, y[v]);
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) y[v][k] = 0.0;
      }
    }
  };
  auto compute = [&](int64_t tile, const int (&g)[NV][8], const double (&y)[NV][8]) {
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      double o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const bool keep = DQMODE == kDqLab ? dq_keep(g[v][k], y[v][k], t) : true;
        const double xr[1] = {keep ? static_cast<double>(g[v][k]) : 0.0};
        o[k] = op.row(acc, consts, xr, keep ? y[v][k] : 0.0, keep, 1);
      }
      if (Op::HAS_OUT && row_out != nullptr) {
        const int64_t row = tile * kTile + (static_cast<int64_t>(v) * kBlock + threadIdx.x) * 8;
#pragma unroll
        for (int k = 0; k < 8; k += 2) stg_stream_d2(row_out + row + k, make_double2(o[k], o[k + 1]));
      }
    }
  };
  int64_t tile = blockIdx.x;
  if (tile < full_tiles) {
    int g[NV][8];
    double y[NV][8];
    load(tile, g, y);
    for (int64_t nxt = tile + gridDim.x; nxt < full_tiles; nxt += gridDim.x) {
      int gn[NV][8];
      double yn[NV][8];
      load(nxt, gn, yn);  // in flight during the arithmetic below
      compute(tile, g, y);
#pragma unroll
      for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          g[v][k] = gn[v][k];
          y[v][k] = yn[v][k];
        }
      tile = nxt;
    }
    compute(tile, g, y);
  }
  if (full_tiles * kTile < n && static_cast<int64_t>(blockIdx.x) == full_tiles % gridDim.x) {  // the ragged end, one row at a time
    ColTable cols;
    cols.ptr[0] = col0;
    cols.i32_mask = 1u;
    DqArgs dq;
    dq.guest = col0;
    dq.price = label;
    dq.t = t;
    for (int64_t r = full_tiles * kTile + threadIdx.x; r < n; r += kBlock)
      process_rows<Op, 1, 1, DQMODE, false>(op, acc, consts, cols, 1, label, dq, r, row_out);
  }
  reduce_and_publish<Op>(acc, dev_consts, partials, ticket, result, ep);
}

}  // namespace dq4ml
