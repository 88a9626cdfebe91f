// wide.hpp -- row-streaming passes over feature tables wider than the 32 columns a register-resident streaming kernel
// holds: prediction / residual / scaled-difference columns with their sums (wide_kernels.cu).
#pragma once

#include <vector>

#include "context.hpp"

namespace dq4ml {

struct WideSums {
  double n, ss_res, sum_y, sum_yy, sum_abs_res, sum_res, sum_res_y, min_res, max_res;
};

// For every row kept by the DQ predicate: pred = x . coef + c0, res = y * y_scale - pred.
// row_out (device, nullable) receives pred (mode 0) or res (mode 1); rows the rules drop get 0.
// With `weight` (instance weights, nullable): rows of weight 0 are skipped, sums.n is the weight sum, sums.ss_res is
// sum w res^2 and row_out carries w * res (mode 1) -- what the weighted loss/gradient pass of a wide table needs.
// mode 2 (HuberAggregator for wide tables, c0 = intercept, y_scale = 1): row_out carries the per-row factor of the
// coefficient gradients, sums.n = weight sum, sums.ss_res = loss sum, sums.sum_y = d/d sigma, sums.sum_res = sum of factors.
// sums are summed over ranks when the ctx is bound.  All data pointers are DEVICE pointers.
int wide_rows_pass(dq4ml_ctx* ctx, const std::vector<const void*>& col_ptrs, const std::vector<int>& dtypes,
                   const double* coef_host, double c0, double y_scale, const double* label, const DeviceDq& dq, int64_t n,
                   int mode, double* row_out, bool reduce_over_ranks, WideSums* sums, const double* weight = nullptr,
                   double sigma = 1.0, double epsilon = 1.35);

// Grow-only device scratch shared by the wide passes (column table, coefficients, Gram partials).
int ensure_wide_scratch(dq4ml_ctx* ctx, size_t bytes);

}  // namespace dq4ml
