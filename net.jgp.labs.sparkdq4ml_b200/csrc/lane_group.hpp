// lane_group.hpp -- the lane-group streaming kernels (lane_group_kernels.cu).
#pragma once

#include <cuda_runtime.h>

#include <cstdint>

#include "kernels.hpp"
#include "stream_reduce.cuh"

namespace dq4ml {

// True when launch_lane_group takes (op, table): the loss/gradient, summarizer, huber, evaluation and gradient-group
// passes over tables of 9..12 columns, the weighted loss/gradient and huber passes over 8..16 features, the weighted
// summarizer at 8 and 16 features -- where they measured ahead of the generic kernel (lane_group_kernels.cu).
bool lane_group_takes(StreamOpKind op, int table_cols);

// Same contract as the generic stream pass (stream_launch.cu): `cols` is the table the host built (for the weighted ops
// its column 0 is the weight), `host_consts` the op's constants (LsqConsts / HuberConsts / EvalConsts or nullptr), and
// the result vector has the layout of the generic kernel's bucket stream_bucket(d).
cudaError_t launch_lane_group(StreamOpKind op, const ColTable& cols, int d, const double* label, const DqArgs& dq,
                              int dq_mode, int64_t n, const void* host_consts, double* row_out, const Scratch& scratch,
                              cudaStream_t stream);

}  // namespace dq4ml
