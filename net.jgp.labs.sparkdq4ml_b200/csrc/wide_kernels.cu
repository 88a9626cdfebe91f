// wide_kernels.cu -- K6 for wide tables: LinearRegressionModel.transform + RegressionMetrics (and the per-row scaled
// difference of LeastSquaresAggregator) when the feature table has more columns than fit in registers.
//
// Spark computes `BLAS.dot(features, coefficients) + intercept` per row (reached from model.transform / summary,
// DataQuality4MachineLearningApp.java:129-139).  Here a thread owns two consecutive rows and walks the columns: every
// column contributes one 128-bit (fp64) or 64-bit (int32) coalesced load per thread, 8 columns in flight, so the pass
// streams the table once at HBM rate; the column pointers and coefficients are warp-uniform L1-resident loads.
#include "device_utils.cuh"
#include "wide.hpp"

namespace dq4ml {
namespace {

struct WideArgs {
  const void* const* col_ptr;
  const unsigned char* col_i32;
  const double* coef;
  int d;
  double c0, y_scale;
  double sigma, epsilon;  // mode 2 (huber)
  const double* label;
  const double* weight;  // instance weights (nullable): the row output and the two leading sums become weighted
  const int* dq_guest;
  const double* dq_price;
  DqThresholds t;
  int has_dq, mode;
  long long n;
  double* row_out;
  double* partials;
  unsigned int* ticket;
  double* result;
};

constexpr int kWideAcc = 9;
__device__ __forceinline__ int wide_kind(int k) { return k == 7 ? kMin : (k == 8 ? kMaxNan : kSum); }

__global__ void __launch_bounds__(kBlock) wide_rows_kernel(const WideArgs a) {
  double acc[kWideAcc] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, __builtin_huge_val(), -__builtin_huge_val()};
  const long long tile_rows = 2LL * kBlock;
  const long long n_tiles = (a.n + tile_rows - 1) / tile_rows;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long row = tile * tile_rows + 2LL * threadIdx.x;
    if (row >= a.n) continue;
    const bool pair = row + 2 <= a.n;
    double d0 = 0.0, d1 = 0.0;
#pragma unroll 8
    for (int j = 0; j < a.d; ++j) {
      const void* base = reinterpret_cast<const void*>(__ldg(reinterpret_cast<const unsigned long long*>(a.col_ptr) + j));
      const double c = __ldg(a.coef + j);
      double x0, x1 = 0.0;
      if (__ldg(a.col_i32 + j)) {
        const int* p = static_cast<const int*>(base) + row;
        if (pair) {
          const int2 v = ldg_stream_i2(p);
          x0 = v.x;
          x1 = v.y;
        } else {
          x0 = __ldg(p);
        }
      } else {
        const double* p = static_cast<const double*>(base) + row;
        if (pair) {
          const double2 v = ldg_stream_d2(p);
          x0 = v.x;
          x1 = v.y;
        } else {
          x0 = __ldg(p);
        }
      }
      d0 = fma(x0, c, d0);
      d1 = fma(x1, c, d1);
    }
    double y0 = 0.0, y1 = 0.0;
    if (a.label != nullptr) {
      if (pair) {
        const double2 v = ldg_stream_d2(a.label + row);
        y0 = v.x;
        y1 = v.y;
      } else {
        y0 = __ldg(a.label + row);
      }
    }
    bool k0 = true, k1 = pair;
    if (a.has_dq) {
      k0 = dq_keep(__ldg(a.dq_guest + row), __ldg(a.dq_price + row), a.t);
      if (pair) k1 = dq_keep(__ldg(a.dq_guest + row + 1), __ldg(a.dq_price + row + 1), a.t);
    }
    // instance weights (weighted LeastSquaresAggregator for wide tables): a row of weight 0 is skipped like a dropped
    // row; a negative or NaN weight poisons the weight sum (Spark: require(weight >= 0.0))
    double w0 = 1.0, w1 = 1.0;
    if (a.weight != nullptr) {
      w0 = __ldg(a.weight + row);
      w1 = pair ? __ldg(a.weight + row + 1) : 0.0;
      if (!(w0 >= 0.0)) w0 = k0 ? __longlong_as_double(0x7ff8000000000000LL) : 0.0;
      if (!(w1 >= 0.0)) w1 = k1 ? __longlong_as_double(0x7ff8000000000000LL) : 0.0;
      k0 = k0 && w0 != 0.0;
      k1 = k1 && w1 != 0.0;
    }
    // dropped rows may carry NaN/Inf garbage: they contribute exact zeros
    const double pred0 = k0 ? d0 + a.c0 : 0.0, pred1 = k1 ? d1 + a.c0 : 0.0;
    const double ys0 = k0 ? y0 * a.y_scale : 0.0, ys1 = k1 ? y1 * a.y_scale : 0.0;
    double r0 = ys0 - pred0, r1 = ys1 - pred1;
    if (a.mode == 2) {
      // HuberAggregator for wide tables: r is the plain residual (label - margin); the row output carries the factor
      // that multiplies x_j / std_j in the gradient, sums: [0] weight sum, [1] loss, [5] sum of factors (d/d intercept),
      // [2] d/d sigma
      auto huber = [&](bool k, double w, double r, double& loss, double& factor, double& dsig) {
        const double rs = r / a.sigma;
        const bool inner = fabs(r) <= a.sigma * a.epsilon;
        loss = k ? w * (inner ? 0.5 * (a.sigma + r * rs) : 0.5 * (a.sigma + 2.0 * a.epsilon * fabs(r) - a.sigma * a.epsilon * a.epsilon)) : 0.0;
        factor = k ? w * (inner ? -rs : (r >= 0.0 ? -a.epsilon : a.epsilon)) : 0.0;
        dsig = k ? w * (inner ? 0.5 * (1.0 - rs * rs) : 0.5 * (1.0 - a.epsilon * a.epsilon)) : 0.0;
      };
      double l0, f0, g0, l1, f1, g1;
      huber(k0, w0, r0, l0, f0, g0);
      huber(k1, w1, r1, l1, f1, g1);
      acc[0] += (k0 ? w0 : 0.0) + (k1 ? w1 : 0.0);
      acc[1] += l0 + l1;
      acc[2] += g0 + g1;
      acc[5] += f0 + f1;
      if (a.row_out != nullptr) {
        if (pair) stg_stream_d2(a.row_out + row, make_double2(f0, f1));
        else a.row_out[row] = f0;
      }
      continue;
    }
    if (a.weight != nullptr) {
      acc[0] += (k0 ? w0 : 0.0) + (k1 ? w1 : 0.0);                 // weightSum
      acc[1] = fma(w0 * r0, r0, fma(w1 * r1, r1, acc[1]));         // sum w r^2
      r0 *= w0;                                                     // the row output carries w r for the gradient pass
      r1 *= w1;
    } else {
      acc[0] += (k0 ? 1.0 : 0.0) + (k1 ? 1.0 : 0.0);
      acc[1] = fma(r0, r0, fma(r1, r1, acc[1]));
    }
    acc[2] += ys0 + ys1;
    acc[3] = fma(ys0, ys0, fma(ys1, ys1, acc[3]));
    acc[4] += fabs(r0) + fabs(r1);
    acc[5] += r0 + r1;
    acc[6] = fma(r0, ys0, fma(r1, ys1, acc[6]));
    acc[7] = combine(kMin, acc[7], combine(kMin, k0 ? r0 : __builtin_huge_val(), k1 ? r1 : __builtin_huge_val()));
    acc[8] = combine(kMaxNan, acc[8], combine(kMaxNan, k0 ? r0 : -__builtin_huge_val(), k1 ? r1 : -__builtin_huge_val()));
    if (a.row_out != nullptr) {
      const double o0 = a.mode == 0 ? pred0 : r0, o1 = a.mode == 0 ? pred1 : r1;
      if (pair) stg_stream_d2(a.row_out + row, make_double2(o0, o1));
      else a.row_out[row] = o0;
    }
  }
  // CTA reduction, then the last CTA to finish reduces the partials in a fixed order (as in stream_reduce_kernel)
  __shared__ double s_red[kBlock / 32];
  __shared__ bool s_last;
  for (int k = 0; k < kWideAcc; ++k) {
    const double v = block_reduce(acc[k], wide_kind(k), s_red);
    if (threadIdx.x == 0) a.partials[static_cast<size_t>(blockIdx.x) * kWideAcc + k] = v;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(a.ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int k = 0; k < kWideAcc; ++k) {
    const int kind = wide_kind(k);
    double v = reduce_identity(kind);
    for (unsigned b = threadIdx.x; b < gridDim.x; b += kBlock) v = combine(kind, v, __ldcg(a.partials + static_cast<size_t>(b) * kWideAcc + k));
    v = block_reduce(v, kind, s_red);
    if (threadIdx.x == 0) a.result[k] = v;
  }
  if (threadIdx.x == 0) *a.ticket = 0u;
}

int cuda_err(cudaError_t e, const char* what) {
  set_error("%s: %s", what, cudaGetErrorString(e));
  return DQ4ML_ERR_CUDA;
}

}  // namespace

int ensure_wide_scratch(dq4ml_ctx* ctx, size_t need) {
  if (need <= ctx->gram_cap) return DQ4ML_OK;
  cudaError_t e;
  if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return cuda_err(e, "sync");
  if (ctx->gram_scratch != nullptr) cudaFree(ctx->gram_scratch);
  ctx->gram_scratch = nullptr;
  ctx->gram_cap = 0;
  if ((e = cudaMalloc(reinterpret_cast<void**>(&ctx->gram_scratch), need)) != cudaSuccess) return cuda_err(e, "wide scratch");
  ctx->gram_cap = need;
  return DQ4ML_OK;
}

int wide_rows_pass(dq4ml_ctx* ctx, const std::vector<const void*>& col_ptrs, const std::vector<int>& dtypes,
                   const double* coef_host, double c0, double y_scale, const double* label, const DeviceDq& dq, int64_t n,
                   int mode, double* row_out, bool reduce_over_ranks, WideSums* sums, const double* weight, double sigma,
                   double epsilon) {
  const int d = static_cast<int>(col_ptrs.size());
  const size_t ptr_b = (static_cast<size_t>(d) * 8 + 255) & ~static_cast<size_t>(255);
  const size_t coef_b = ptr_b;
  const size_t flag_b = (static_cast<size_t>(d) + 255) & ~static_cast<size_t>(255);
  int rc = ensure_wide_scratch(ctx, ptr_b + coef_b + flag_b + 256);
  if (rc != DQ4ML_OK) return rc;
  char* base = ctx->gram_scratch;
  std::vector<unsigned char> flags(d > 0 ? d : 1);
  for (int j = 0; j < d; ++j) flags[j] = dtypes[j] == DQ4ML_I32 ? 1 : 0;
  cudaError_t e;
  if (d > 0) {
    if ((e = cudaMemcpyAsync(base, col_ptrs.data(), static_cast<size_t>(d) * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return cuda_err(e, "table upload");
    if ((e = cudaMemcpyAsync(base + ptr_b, coef_host, static_cast<size_t>(d) * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return cuda_err(e, "coefficient upload");
    if ((e = cudaMemcpyAsync(base + ptr_b + coef_b, flags.data(), static_cast<size_t>(d), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return cuda_err(e, "table upload");
    if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return cuda_err(e, "sync");  // host staging vectors die with this frame
  }
  WideArgs a;
  a.col_ptr = reinterpret_cast<const void* const*>(base);
  a.coef = reinterpret_cast<const double*>(base + ptr_b);
  a.col_i32 = reinterpret_cast<const unsigned char*>(base + ptr_b + coef_b);
  a.d = d;
  a.c0 = c0;
  a.y_scale = y_scale;
  a.sigma = sigma;
  a.epsilon = epsilon;
  a.label = label;
  a.weight = weight;
  a.dq_guest = dq.guest;
  a.dq_price = dq.price;
  a.t.min_price = dq.params.min_price;
  a.t.max_price = dq.params.max_price;
  a.t.min_guest = dq.params.min_guest;
  a.has_dq = dq.guest != nullptr ? 1 : 0;
  a.mode = mode;
  a.n = n;
  a.row_out = row_out;
  a.partials = ctx->scratch.partials;
  a.ticket = ctx->scratch.ticket;
  a.result = ctx->scratch.result;
  const int64_t n_tiles = (n + 2 * kBlock - 1) / (2 * kBlock);
  int64_t grid = 4LL * ctx->sm_count;
  if (grid > n_tiles) grid = n_tiles;
  if (grid < 1) grid = 1;
  if (ctx->pass_timing) cudaEventRecord(ctx->pass_ev0, ctx->stream);
  wide_rows_kernel<<<static_cast<unsigned>(grid), kBlock, 0, ctx->stream>>>(a);
  if ((e = cudaGetLastError()) != cudaSuccess) return cuda_err(e, "wide rows kernel");
  ++ctx->launches;
  if (ctx->pass_timing) {
    cudaEventRecord(ctx->pass_ev1, ctx->stream);
    ctx->pass_event_pending = true;
  }
  if (reduce_over_ranks && ctx->nccl != nullptr) {
    rc = nccl_allreduce_device(ctx, ctx->scratch.result, kWideAcc);
    if (rc != DQ4ML_OK) return rc;
  }
  if ((e = cudaMemcpyAsync(ctx->h_result, ctx->scratch.result, sizeof(double) * kWideAcc, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) return cuda_err(e, "D2H");
  if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return cuda_err(e, "wide rows pass");
  if (ctx->pass_event_pending) {
    ctx->pass_event_pending = false;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ctx->pass_ev0, ctx->pass_ev1) == cudaSuccess) {
      ctx->pass_ms_total += ms;
      ++ctx->pass_count;
    }
  }
  if (reduce_over_ranks && ctx->nccl == nullptr && ctx->allreduce_cb != nullptr) {
    if (ctx->allreduce_cb(ctx->allreduce_user, ctx->h_result, kWideAcc) != 0) {
      set_error("all-reduce callback failed");
      return DQ4ML_ERR_COMM;
    }
  }
  if (sums != nullptr) {
    sums->n = ctx->h_result[0];
    sums->ss_res = ctx->h_result[1];
    sums->sum_y = ctx->h_result[2];
    sums->sum_yy = ctx->h_result[3];
    sums->sum_abs_res = ctx->h_result[4];
    sums->sum_res = ctx->h_result[5];
    sums->sum_res_y = ctx->h_result[6];
    sums->min_res = ctx->h_result[7];
    sums->max_res = ctx->h_result[8];
  }
  return DQ4ML_OK;
}

}  // namespace dq4ml
