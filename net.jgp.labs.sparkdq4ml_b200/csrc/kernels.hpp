// kernels.hpp -- host-callable launchers of the sm_100a kernels.  All pointers are DEVICE pointers here; the C ABI
// layer (api.cpp) resolves host buffers before calling in.  Every launcher enqueues on `stream`, returns the CUDA
// status of the launch, and adds the number of kernels it launched to *launches.
#pragma once

#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/dq4ml.h"

namespace dq4ml {

constexpr int kMaxCols = 32;  // widest column group handled by one streaming launch

struct DeviceCols {
  const void* ptr[kMaxCols];
  int dtype[kMaxCols];
  int n = 0;
};

struct DeviceDq {
  const int32_t* guest = nullptr;  // nullptr = no fused DQ
  const double* price = nullptr;
  dq4ml_dq_params params{};
};

// How the last CTA of a streaming pass hands its result vector over (epilogue.cuh): to mapped pinned host memory, after
// an exchange through the peer mailboxes when several ranks are bound.
constexpr int kXrMaxRanks = 16;
constexpr int kXrSlotDoubles = 128;  // longest published vector: SummCoreOp<32> = 67 sums + 33 centres
constexpr unsigned kXrErrorTag = 0xffffffffu;  // tag of word pair 0 of the host buffer when the exchange timed out
struct XrEpilogue {
  unsigned long long* host_out = nullptr;  // device alias of the pinned result buffer (tagged word pairs, epilogue.cuh);
                                           // nullptr = leave the vector in scratch.result
  unsigned long long host_seq = 0;         // tag (low 32 bits) the words of this pass carry in the host buffer
  int world = 1, rank = 0;                 // world > 1: exchange through the mailboxes below
  int gather = 0;                          // 0: fold the ranks' vectors in rank order; 1: hand the host all of them
  int extra_consts = 0;                    // append this many of the pass's device constants to the vector
  unsigned long long xr_seq = 0;           // number of this cross-rank pass (identical on every rank)
  unsigned char* peer[kXrMaxRanks] = {};   // rank r's mailbox as mapped into this process; peer[rank] is our own
};

// Scratch a ctx lends to the launchers.
struct Scratch {
  double* partials;        // grid x NACC doubles
  size_t partials_bytes;
  unsigned int* ticket;    // zero between launches
  double* result;          // >= 4096 doubles
  double* consts;          // >= 64 doubles
  int sm_count;
  XrEpilogue ep;           // armed per pass by api.cpp
};

enum StreamOpKind {
  kOpWls = 0, kOpLsq = 1, kOpSummCore = 2, kOpSummFull = 3, kOpEval = 4, kOpGrad = 5,
  // weighted variants: column 0 of the table is the instance weight, the features follow (so at most kMaxCols - 1 of them)
  kOpSummCoreW = 6, kOpLsqW = 7, kOpWResid = 8,
  // huber loss (LinearRegression.setLoss("huber")): HuberAggregator, plain and with the weight as column 0
  kOpHuber = 9, kOpHuberW = 10
};

struct LsqConsts {
  double eff[kMaxCols];
  double offset;
  double inv_ystd;
};
struct HuberConsts {
  double eff[kMaxCols];  // coefficient / featuresStd per column of the streamed table (0 for the weight column)
  double intercept, sigma, epsilon;
};
struct EvalConsts {
  double coef[kMaxCols];
  double intercept;
  int mode;  // 0: per-row output = prediction, 1: residual
};

// Number of accumulators the op produces in Scratch::result for a column-group bucket (1,2,4,8,16,32).
int stream_bucket(int d);
int stream_nacc(StreamOpKind op, int bucket);

// The generic streaming pass.  `host_consts` is LsqConsts / EvalConsts for those ops, ignored otherwise; SummCore /
// SummFull read their centres from scratch.consts (device).  row_out may be nullptr.
cudaError_t launch_stream_pass(StreamOpKind op, const DeviceCols& cols, const double* label, const DeviceDq& dq,
                               int64_t n, const void* host_consts, double* row_out, const Scratch& scratch,
                               cudaStream_t stream, int64_t* launches);

// centres[j] = mean of the first min(n, 256) values of column j: the provisional centre of the shifted-sum summarizer.
// Writes label_slot + 1 doubles to scratch.consts: features at [0, cols.n), zeros up to label_slot, the label's centre
// at label_slot (0 when label is null).
cudaError_t launch_column_heads(const DeviceCols& cols, const double* label, int label_slot, int64_t n,
                                const Scratch& scratch, cudaStream_t stream, int64_t* launches);

// DQ UDF columns.  err_flag (device int, pre-zeroed) is set to 1 if a null reaches minimumPriceRule.
cudaError_t launch_udf_min_price(const double* price, const uint8_t* valid, int64_t n, double* out,
                                 const dq4ml_dq_params& p, int* err_flag, cudaStream_t stream, int64_t* launches);
cudaError_t launch_udf_price_corr(const double* price, const int32_t* guest, const uint8_t* pvalid,
                                  const uint8_t* gvalid, int64_t n, double* out, const dq4ml_dq_params& p,
                                  cudaStream_t stream, int64_t* launches);

// Stable filter-compaction (single pass, decoupled look-back).  tile_state: dq_filter_state_bytes(n) bytes of scratch
// that the launcher zeroes itself; counters: 2 x uint64 device ([0] survivors, [1] rows after rule 1).
constexpr int64_t kFilterStateTile = 2048;  // rows per status word (the tile of the persistent filter kernel)
size_t dq_filter_state_bytes(int64_t n);
cudaError_t launch_dq_filter(const int32_t* guest, const double* price, int64_t n, const dq4ml_dq_params& p,
                             int32_t* guest_out, double* price_out, unsigned long long* tile_state,
                             unsigned long long* counters, cudaStream_t stream, int64_t* launches);

cudaError_t launch_synth(const dq4ml_synth_params& p, int64_t row0, int64_t n, int32_t* guest, double* price,
                         double* extra, int64_t extra_stride, cudaStream_t stream, int64_t* launches);
cudaError_t launch_rows_digest(const int32_t* guest, const double* price, int64_t n, uint64_t seed,
                               unsigned long long* out2 /* device, zeroed by the launcher */, cudaStream_t stream,
                               int64_t* launches);

}  // namespace dq4ml
