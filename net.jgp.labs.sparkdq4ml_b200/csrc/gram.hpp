// gram.hpp -- WLS moments for wide feature tables (d > 8): the X^T X / X^T y Gram pass (gram_kernels.cu).
#pragma once

#include <vector>

#include "context.hpp"

namespace dq4ml {

// Same output layout as dq4ml_wls_moments.  col_ptrs / label / weight (nullable: all 1) / dq are DEVICE pointers,
// 32-byte aligned.
int gram_wls_moments(dq4ml_ctx* ctx, const std::vector<const void*>& col_ptrs, const std::vector<int>& dtypes,
                     const double* label, const double* weight, const DeviceDq& dq, int64_t n, double* out);

}  // namespace dq4ml
