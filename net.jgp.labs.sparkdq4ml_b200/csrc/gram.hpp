// gram.hpp -- WLS moments for wide feature tables (d > 8): the X^T X / X^T y Gram pass (gram_kernels.cu).
#pragma once

#include <vector>

#include "context.hpp"

namespace dq4ml {

// Same output layout as dq4ml_wls_moments.  col_ptrs / label / weight (nullable: all 1) / dq are DEVICE pointers,
// 32-byte aligned.
int gram_wls_moments(dq4ml_ctx* ctx, const std::vector<const void*>& col_ptrs, const std::vector<int>& dtypes,
                     const double* label, const double* weight, const DeviceDq& dq, int64_t n, double* out);

// The tcgen05 / TMEM tier of the same pass (gram_tf32_kernels.cu): 3xTF32 split products, fp32 accumulation in TMEM over
// bounded row spans, fp64 across spans.  Same arguments and output layout; 1 <= d <= 256; relaxed tolerance (1e-5).
bool gram_tf32_supported(int d);
int gram_tf32_wls_moments(dq4ml_ctx* ctx, const std::vector<const void*>& col_ptrs, const std::vector<int>& dtypes,
                          const double* label, const double* weight, const DeviceDq& dq, int64_t n, double* out);

}  // namespace dq4ml
