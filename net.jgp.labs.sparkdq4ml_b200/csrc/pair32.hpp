// pair32.hpp -- launcher of the lane-pair kernels for the 17..32-column summarizer and loss/gradient passes
// (pair32_kernels.cu).  Included from .cu files only (ColTable / DqArgs are device-side types).
#pragma once

#include "kernels.hpp"
#include "stream_reduce.cuh"

namespace dq4ml {

constexpr int kPairLsq = 0, kPairSumm = 1;

// Same inputs, scratch and result layout as launch_stream_pass(kOpLsq / kOpSummCore) with the 32-column bucket; `lsq`
// is read for kPairLsq only, the summarizer's centres come from scratch.consts[0..32] (label at 32).
cudaError_t launch_pair32(int op, const ColTable& cols, int d, const double* label, const DqArgs& dq, int dq_mode, int64_t n,
                          const LsqConsts* lsq, const Scratch& scratch, cudaStream_t stream);

}  // namespace dq4ml
