// stream_launch.cu -- instantiates stream_reduce_kernel for every (op, column-group bucket, DQ mode) the path uses
// and picks one at run time.  Buckets 1,2 stream 8 rows per thread per tile, buckets 4,8,16 stream 4, bucket 32 streams 2
// (register budget); grids are sized to (CTAs resident per SM) x (SM count) so the whole pass is one wave of persistent CTAs.
// Dispatch of a pass by (op, table columns), each choice from profiles/r2_bucket_sweep*.jsonl: the lane-group kernels
// (lane_group_kernels.cu) take 9..12 columns, the weighted loss/gradient and huber passes over 8..16 features, and the
// evaluation / huber / gradient-group passes over 17..32 columns; the lane-pair / lane-quad kernels (pair32_kernels.cu)
// take the summarizer and the loss/gradient pass over 17..32 columns; everything else is the generic kernel here.
#include <cstdlib>
#include <map>
#include <mutex>
#include <type_traits>

#include "kernels.hpp"
#include "ops.cuh"
#include "lane_group.hpp"
#include "pair32.hpp"
#include "stream_reduce.cuh"

namespace dq4ml {

int stream_bucket(int d) {
  int b = 1;
  while (b < d) b <<= 1;
  return b;
}

int stream_nacc(StreamOpKind op, int bucket) {
  switch (op) {
    case kOpWls: return 3 + 2 * bucket + bucket * (bucket + 1) / 2;
    case kOpLsq: return 2 + bucket;
    case kOpSummCore: return 1 + 2 * (bucket + 1);
    case kOpSummFull: return 1 + 7 * bucket;
    case kOpEval: return 9;
    case kOpGrad: return bucket;
    case kOpSummCoreW: return 3 + 2 * bucket;
    case kOpLsqW: return 2 + bucket;
    case kOpWResid: return 3;
    case kOpHuber: return 4 + bucket;
    case kOpHuberW: return 4 + bucket;
  }
  return 0;
}

namespace {

struct PassArgs {
  ColTable cols;
  int d;
  const double* label;
  DqArgs dq;
  int dq_mode;
  int64_t n;
  const double* dev_consts;
  double* row_out;
  Scratch scratch;
  cudaStream_t stream;
};

// A table whose shape is known at compile time (every column of the bucket present, at most one int32 column, in
// position 0 or 1): the loss/gradient, summarizer, huber, evaluation and gradient-group passes then run with a load
// phase without per-column tests (stream_reduce.cuh, DENSE).  Returns 0, or 1 + the bit mask of the int32 columns for a
// full bucket, 5 + the mask for a narrower table of the same column types.
inline int dense_id(const PassArgs& a, int dmax) {
  static const char* off = std::getenv("DQ4ML_STREAM_DENSE");  // DQ4ML_STREAM_DENSE=0: the general load phase (A/B)
  if (off != nullptr && off[0] == '0') return 0;
  if (a.d > dmax || a.cols.i32_mask > 2u) return 0;
  return (a.d == dmax ? 1 : 5) + static_cast<int>(a.cols.i32_mask);
}
// ops whose narrower tables also gain from compile-time column types (the width test stays): the dot-product ops
template <class Op, int DMAX>
constexpr bool kPartialOp = std::is_same<Op, LsqOp<DMAX>>::value || std::is_same<Op, EvalOp<DMAX>>::value ||
                            std::is_same<Op, HuberOp<DMAX>>::value || std::is_same<Op, GradOp<DMAX>>::value;
template <class Op, int DMAX>
constexpr bool kPartialWeightedOp = std::is_same<Op, LsqWOp<DMAX>>::value || std::is_same<Op, HuberWOp<DMAX>>::value;
template <class Op, int DMAX>
constexpr bool kDenseOp = std::is_same<Op, LsqOp<DMAX>>::value || std::is_same<Op, SummCoreOp<DMAX>>::value ||
                          std::is_same<Op, EvalOp<DMAX>>::value || std::is_same<Op, HuberOp<DMAX>>::value ||
                          std::is_same<Op, GradOp<DMAX>>::value;
template <class Op, int DMAX>
constexpr bool kDenseWeightedOp = std::is_same<Op, LsqWOp<DMAX>>::value || std::is_same<Op, SummCoreWOp<DMAX>>::value ||
                                  std::is_same<Op, HuberWOp<DMAX>>::value;

template <class Op, int DMAX, int ROWS, int DQMODE, int DENSE = 0>
cudaError_t launch_one(const Op& op, const PassArgs& a) {
  auto kernel = stream_reduce_kernel<Op, DMAX, ROWS, DQMODE, DENSE>;
  static std::mutex mu;
  static std::map<int, int> blocks_per_sm;  // per device
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  int per_sm = 0;
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = blocks_per_sm.find(dev);
    if (it == blocks_per_sm.end()) {
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kBlock, 0);
      if (e != cudaSuccess) return e;
      if (per_sm < 1) per_sm = 1;
      // measured (tools/k2_variants.cu, 10^8 rows of 12 B): 4 CTAs/SM stream at 6.0-6.2 TB/s, 6 CTAs/SM at 4.9-5.7
      if (ROWS == 8 && per_sm > 4) per_sm = 4;  // measured on the bench workload: 3 -> 5.76, 4 -> 6.36, 6 -> 5.45 TB/s
      if (const char* env = getenv("DQ4ML_STREAM_CTAS")) { const int v = atoi(env); if (v >= 1) per_sm = v; }
      blocks_per_sm[dev] = per_sm;
    } else {
      per_sm = it->second;
    }
  }
  const int64_t tile_rows = static_cast<int64_t>(kBlock) * ROWS;
  const int64_t n_tiles = (a.n + tile_rows - 1) / tile_rows;  // the partial tile is taken by CTA (full_tiles % grid) < grid
  int64_t grid = static_cast<int64_t>(per_sm) * a.scratch.sm_count;
  if (grid > n_tiles) grid = n_tiles;
  if (grid < 1) grid = 1;
  const size_t need = static_cast<size_t>(grid) * Op::NACC * sizeof(double);
  if (need > a.scratch.partials_bytes) return cudaErrorMemoryAllocation;
  kernel<<<static_cast<unsigned>(grid), kBlock, 0, a.stream>>>(op, a.cols, a.d, a.label, a.dq, a.n, a.dev_consts,
                                                                a.row_out, a.scratch.partials, a.scratch.ticket,
                                                                a.scratch.result, a.scratch.ep);
  return cudaGetLastError();
}

// The software-pipelined kernel for the lab's own shape: one int32 feature column + the label, DQ in the lab layout or off
// (stream_pipe_kernel: 24 rows per thread in the arithmetic, 24 more in flight, one CTA per SM).
constexpr int kPipeVectors = 3;
template <class Op, int DQMODE>
cudaError_t launch_pipe(const Op& op, const PassArgs& a) {
  constexpr int64_t kTile = static_cast<int64_t>(kBlock) * 8 * kPipeVectors;
  const int64_t n_tiles = (a.n + kTile - 1) / kTile;
  int64_t grid = a.scratch.sm_count;
  if (grid > n_tiles) grid = n_tiles;
  if (grid < 1) grid = 1;
  if (static_cast<size_t>(grid) * Op::NACC * sizeof(double) > a.scratch.partials_bytes) return cudaErrorMemoryAllocation;
  stream_pipe_kernel<Op, kPipeVectors, DQMODE><<<static_cast<unsigned>(grid), kBlock, 0, a.stream>>>(
      op, static_cast<const int*>(a.cols.ptr[0]), a.label, a.dq.t, a.n, a.dev_consts, a.row_out, a.scratch.partials, a.scratch.ticket,
      a.scratch.result, a.scratch.ep);
  return cudaGetLastError();
}
inline bool pipe_takes(const PassArgs& a) {
  static const char* off = std::getenv("DQ4ML_STREAM_PIPE");  // DQ4ML_STREAM_PIPE=0: the grid-strided kernel (A/B)
  if (off != nullptr && off[0] == '0') return false;
  return a.d == 1 && (a.cols.i32_mask & 1u) != 0u && (a.dq_mode == kDqNone || a.dq_mode == kDqLab);
}

template <class Op, int DMAX, int ROWS>
cudaError_t launch_mode(const Op& op, const PassArgs& a) {
  if constexpr (DMAX == 1) {
    if (pipe_takes(a)) return a.dq_mode == kDqLab ? launch_pipe<Op, kDqLab>(op, a) : launch_pipe<Op, kDqNone>(op, a);
  }
  if constexpr ((DMAX == 32 || DMAX == 16 || DMAX == 8) && kDenseOp<Op, DMAX>) {
    const int dense = dense_id(a, DMAX);
    if (dense == 2) {  // column 0 int32: the lab layout, or the same table without the rules
      if (a.dq_mode == kDqLab) return launch_one<Op, DMAX, ROWS, kDqLab, 2>(op, a);
      if (a.dq_mode == kDqNone) return launch_one<Op, DMAX, ROWS, kDqNone, 2>(op, a);
    } else if (dense == 1 && a.dq_mode == kDqNone) {
      return launch_one<Op, DMAX, ROWS, kDqNone, 1>(op, a);
    }
    if constexpr ((DMAX == 32 || DMAX == 16) && kPartialOp<Op, DMAX>) {
      if (dense == 6) {
        if (a.dq_mode == kDqLab) return launch_one<Op, DMAX, ROWS, kDqLab, 6>(op, a);
        if (a.dq_mode == kDqNone) return launch_one<Op, DMAX, ROWS, kDqNone, 6>(op, a);
      } else if (dense == 5 && a.dq_mode == kDqNone) {
        return launch_one<Op, DMAX, ROWS, kDqNone, 5>(op, a);
      }
    }
  }
  switch (a.dq_mode) {
    case kDqNone: return launch_one<Op, DMAX, ROWS, kDqNone>(op, a);
    case kDqSeparate: return launch_one<Op, DMAX, ROWS, kDqSeparate>(op, a);
    default: return launch_one<Op, DMAX, ROWS, kDqLab>(op, a);
  }
}

// the weighted ops never see the lab layout (column 0 is the weight), so two DQ modes are enough
template <class Op, int DMAX, int ROWS>
cudaError_t launch_mode2(const Op& op, const PassArgs& a) {
  if constexpr ((DMAX == 8 || DMAX == 16 || DMAX == 32) && kDenseWeightedOp<Op, DMAX>) {  // weight, then all fp64 or guest (int32) first
    const int dense = dense_id(a, DMAX);
    if (dense == 1) return a.dq_mode == kDqNone ? launch_one<Op, DMAX, ROWS, kDqNone, 1>(op, a) : launch_one<Op, DMAX, ROWS, kDqSeparate, 1>(op, a);
    if (dense == 3) return a.dq_mode == kDqNone ? launch_one<Op, DMAX, ROWS, kDqNone, 3>(op, a) : launch_one<Op, DMAX, ROWS, kDqSeparate, 3>(op, a);
    if constexpr ((DMAX == 16 || DMAX == 32) && kPartialWeightedOp<Op, DMAX>) {
      if (dense == 5) return a.dq_mode == kDqNone ? launch_one<Op, DMAX, ROWS, kDqNone, 5>(op, a) : launch_one<Op, DMAX, ROWS, kDqSeparate, 5>(op, a);
      if (dense == 7) return a.dq_mode == kDqNone ? launch_one<Op, DMAX, ROWS, kDqNone, 7>(op, a) : launch_one<Op, DMAX, ROWS, kDqSeparate, 7>(op, a);
    }
  }
  if (a.dq_mode == kDqNone) return launch_one<Op, DMAX, ROWS, kDqNone>(op, a);
  return launch_one<Op, DMAX, ROWS, kDqSeparate>(op, a);
}

template <class L, int DMAX>
L make_lsq(const LsqConsts& c) {
  L op;
  for (int j = 0; j < DMAX; ++j) op.eff[j] = c.eff[j];
  op.offset = c.offset;
  op.inv_ystd = c.inv_ystd;
  return op;
}
template <class H, int DMAX>
H make_huber(const HuberConsts& c) {
  H op;
  for (int j = 0; j < DMAX; ++j) op.eff[j] = c.eff[j];
  op.intercept = c.intercept;
  op.sigma = c.sigma;
  op.epsilon = c.epsilon;
  return op;
}
template <int DMAX>
EvalOp<DMAX> make_eval(const EvalConsts& c) {
  EvalOp<DMAX> op;
  for (int j = 0; j < DMAX; ++j) op.coef[j] = c.coef[j];
  op.intercept = c.intercept;
  op.mode = c.mode;
  return op;
}

template <int DMAX, int ROWS>
cudaError_t launch_bucket_rows(StreamOpKind kind, const PassArgs& a, const void* host_consts) {
  switch (kind) {
    case kOpWls:
      if constexpr (DMAX <= 8) return launch_mode<WlsOp<DMAX>, DMAX, ROWS>(WlsOp<DMAX>{}, a);
      return cudaErrorInvalidValue;
    case kOpLsq: return launch_mode<LsqOp<DMAX>, DMAX, ROWS>(make_lsq<LsqOp<DMAX>, DMAX>(*static_cast<const LsqConsts*>(host_consts)), a);
    case kOpSummCore: return launch_mode<SummCoreOp<DMAX>, DMAX, ROWS>(SummCoreOp<DMAX>{}, a);
    case kOpSummFull:
      if constexpr (DMAX <= 4) return launch_mode<SummFullOp<DMAX>, DMAX, ROWS>(SummFullOp<DMAX>{}, a);
      return cudaErrorInvalidValue;
    case kOpEval: return launch_mode<EvalOp<DMAX>, DMAX, ROWS>(make_eval<DMAX>(*static_cast<const EvalConsts*>(host_consts)), a);
    case kOpGrad: return launch_mode<GradOp<DMAX>, DMAX, ROWS>(GradOp<DMAX>{}, a);  // always kDqNone
    case kOpSummCoreW:
      if constexpr (DMAX >= 2) return launch_mode2<SummCoreWOp<DMAX>, DMAX, ROWS>(SummCoreWOp<DMAX>{}, a);
      return cudaErrorInvalidValue;
    case kOpLsqW:
      if constexpr (DMAX >= 2) return launch_mode2<LsqWOp<DMAX>, DMAX, ROWS>(make_lsq<LsqWOp<DMAX>, DMAX>(*static_cast<const LsqConsts*>(host_consts)), a);
      return cudaErrorInvalidValue;
    case kOpWResid:
      if constexpr (DMAX == 2) return launch_mode2<WResidOp<DMAX>, DMAX, ROWS>(WResidOp<DMAX>{}, a);
      return cudaErrorInvalidValue;
    case kOpHuber: return launch_mode<HuberOp<DMAX>, DMAX, ROWS>(make_huber<HuberOp<DMAX>, DMAX>(*static_cast<const HuberConsts*>(host_consts)), a);
    case kOpHuberW:
      if constexpr (DMAX >= 2) return launch_mode2<HuberWOp<DMAX>, DMAX, ROWS>(make_huber<HuberWOp<DMAX>, DMAX>(*static_cast<const HuberConsts*>(host_consts)), a);
      return cudaErrorInvalidValue;
  }
  return cudaErrorInvalidValue;
}

template <int DMAX>
cudaError_t launch_bucket(StreamOpKind kind, const PassArgs& a, const void* host_consts) {
  // rows per thread per tile: what keeps ~100 bytes per thread in flight without spilling (tools/k2_variants.cu: 8 rows
  // with 256-bit loads at 3 CTAs/SM stream a 12 B/row table at 6.2 TB/s, 4 rows at 6 CTAs/SM at 5.8)
  // The 16-column bucket also takes 4 rows (256-bit loads for f64) although that leaves one CTA of ~220 registers per
  // SM: 10^9 rows x 132 B/row per pass went from 0.88 to 0.97 of the measured HBM peak (22.8 -> 20.9 ms).
  constexpr int ROWS = DMAX <= 2 ? 8 : (DMAX <= 16 ? 4 : 2);
  return launch_bucket_rows<DMAX, ROWS>(kind, a, host_consts);
}

}  // namespace

cudaError_t launch_stream_pass(StreamOpKind op, const DeviceCols& cols, const double* label, const DeviceDq& dq,
                               int64_t n, const void* host_consts, double* row_out, const Scratch& scratch,
                               cudaStream_t stream, int64_t* launches) {
  if (cols.n < 0 || cols.n > kMaxCols) return cudaErrorInvalidValue;
  PassArgs a;
  a.d = cols.n;
  a.cols.i32_mask = 0;
  for (int j = 0; j < kMaxStreamD; ++j) a.cols.ptr[j] = nullptr;
  for (int j = 0; j < cols.n; ++j) {
    a.cols.ptr[j] = cols.ptr[j];
    if (cols.dtype[j] == DQ4ML_I32) a.cols.i32_mask |= 1u << j;
  }
  a.label = label;
  a.dq.guest = dq.guest;
  a.dq.price = dq.price;
  a.dq.t.min_price = dq.params.min_price;
  a.dq.t.max_price = dq.params.max_price;
  a.dq.t.min_guest = dq.params.min_guest;
  if (dq.guest == nullptr) {
    a.dq_mode = kDqNone;
  } else if ((op < kOpSummCoreW || op == kOpHuber) && cols.n > 0 && cols.dtype[0] == DQ4ML_I32 && cols.ptr[0] == dq.guest &&
             label != nullptr && label == dq.price) {
    a.dq_mode = kDqLab;
  } else {
    a.dq_mode = kDqSeparate;
  }
  a.n = n;
  a.dev_consts = scratch.consts;
  a.row_out = row_out;
  a.scratch = scratch;
  a.stream = stream;
  cudaError_t e;
  // Which tables the generic kernel takes with a compile-time load phase (at most one int32 column, in position 0, or 1
  // behind a weight): a full bucket of 8, 16 or 32 columns for every op (1.07-1.13 of the measured copy peak, ahead of
  // the lane-group / lane-pair kernels); narrower tables for the dot-product ops where they measured ahead
  // (profiles/r2_bucket_sweep_dense.jsonl) -- 9..15 and 21..31 columns for the loss/gradient pass (17..20 stay lane
  // pairs), 9..31 for evaluation, huber and the gradient groups, weighted loss/gradient and huber over 9..14 and 17..30
  // features.  The summarizer's narrower tables stay on the lane-group / lane-pair kernels.
  const bool weighted_op = op == kOpLsqW || op == kOpSummCoreW || op == kOpHuberW;
  const bool dense_family = weighted_op || op == kOpLsq || op == kOpSummCore || op == kOpEval || op == kOpHuber || op == kOpGrad;
  const bool dense_mode = weighted_op ? a.dq_mode != kDqLab : a.dq_mode != kDqSeparate;
  const bool dense_mask = weighted_op ? (a.cols.i32_mask == 0u || a.cols.i32_mask == 2u) : a.cols.i32_mask <= 1u;
  const bool dense_ok = dense_family && dense_mode && dense_mask && dense_id(a, stream_bucket(cols.n > 0 ? cols.n : 1)) != 0 &&
                        !(a.dq_mode == kDqLab && a.cols.i32_mask != 1u);
  bool full_bucket = dense_ok && (cols.n == 8 || cols.n == 16 || cols.n == 32);
  if (dense_ok && !full_bucket) {
    const int w = cols.n;
    if (op == kOpLsq) full_bucket = (w >= 9 && w <= 15) || (w >= 21 && w <= 31);
    else if (op == kOpEval || op == kOpHuber || op == kOpGrad) full_bucket = w >= 9 && w <= 31;
    else if (op == kOpLsqW || op == kOpHuberW) full_bucket = (w >= 10 && w <= 15) || (w >= 18 && w <= 31);
  }
  if (!full_bucket && lane_group_takes(op, cols.n)) {  // 9..12 columns, 8..16 weighted features: the lane-group kernels (lane_group_kernels.cu)
    e = launch_lane_group(op, a.cols, a.d, a.label, a.dq, a.dq_mode, a.n, host_consts, row_out, scratch, stream);
    if (e == cudaSuccess) ++*launches;
    return e;
  }
  if (!full_bucket && (op == kOpLsq || op == kOpSummCore) && cols.n > 16) {  // 17..32 columns: the lane-pair / lane-quad kernels (pair32_kernels.cu)
    e = launch_pair32(op == kOpLsq ? kPairLsq : kPairSumm, a.cols, a.d, a.label, a.dq, a.dq_mode, a.n,
                      static_cast<const LsqConsts*>(host_consts), scratch, stream);
    if (e == cudaSuccess) ++*launches;
    return e;
  }
  switch (stream_bucket(cols.n > 0 ? cols.n : 1)) {
    case 1: e = launch_bucket<1>(op, a, host_consts); break;
    case 2: e = launch_bucket<2>(op, a, host_consts); break;
    case 4: e = launch_bucket<4>(op, a, host_consts); break;
    case 8: e = launch_bucket<8>(op, a, host_consts); break;
    case 16: e = launch_bucket<16>(op, a, host_consts); break;
    default: e = launch_bucket<32>(op, a, host_consts); break;
  }
  if (e == cudaSuccess) ++*launches;
  return e;
}

}  // namespace dq4ml
