// gram.hpp -- WLS moments for wide feature tables (d > 8): the X^T X / X^T y Gram pass.
#pragma once

#include "context.hpp"

namespace dq4ml {

// Same output layout as dq4ml_wls_moments.  All column pointers are device pointers.
int gram_wls_moments(dq4ml_ctx* ctx, const DeviceCols& cols, const double* label, const DeviceDq& dq, int64_t n,
                     double* out);

}  // namespace dq4ml
