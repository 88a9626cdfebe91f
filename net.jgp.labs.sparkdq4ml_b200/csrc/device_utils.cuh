// device_utils.cuh -- shared device helpers: streaming 128-bit loads, the DQ predicate, block reductions.
#pragma once

#include <cuda_runtime.h>

#include <cstdint>

namespace dq4ml {

constexpr int kBlock = 256;        // threads per CTA of the streaming-reduce kernels
constexpr int kMaxStreamD = 32;    // widest column group a streaming kernel keeps in registers

// Thresholds of the two rules (MinimumPriceDataQualityService.java:5-9, PriceCorrelationDataQualityService.java:6).
struct DqThresholds {
  double min_price;
  double max_price;
  int min_guest;
};

// Both rules + both filters for one row (DataQuality4MachineLearningApp.java:68-90):
//   r1 = price < min_price ? -1 : price        keep iff r1 > 0   (Spark SQL ordering: NaN > 0 is true)
//   r2 = guest < min_guest && r1 > max_price ? -1 : r1           keep iff r2 > 0
// Java's `<`/`>` are IEEE (false on NaN), so a NaN price passes through both rules and both filters.
// Reduced to three fp64 compares (the fp64 pipe is shared with the accumulations of a 12 B/row pass):
//   r1 > 0 (Spark ordering)  <=>  !(price < min_price) && !(price <= 0)   [NaN fails both tests and is kept]
//   r2 > 0                   <=>  r1 > 0 && !(guest < min_guest && price > max_price)
__device__ __forceinline__ bool dq_keep_rule1(double price, const DqThresholds& t) {
  return !(price < t.min_price) && !(price <= 0.0);
}
__device__ __forceinline__ bool dq_keep(int guest, double price, const DqThresholds& t) {
  return dq_keep_rule1(price, t) && !(guest < t.min_guest && price > t.max_price);
}

// Streaming loads: read-only path, do not allocate in L1 (every byte is touched once per pass).
__device__ __forceinline__ int4 ldg_stream_i4(const void* p) {
  int4 r;
  asm("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ int2 ldg_stream_i2(const void* p) {
  int2 r;
  asm("ld.global.nc.L1::no_allocate.v2.s32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ double2 ldg_stream_d2(const void* p) {
  double2 r;
  asm("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_stream_d2(void* p, double2 v) {
  asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}

// ROWS consecutive values of one column starting at `row`, widened to double (VectorAssembler's cast for int32
// columns).  VEC: the column base is 16-byte aligned and row % ROWS == 0, so 128-bit / 64-bit vector loads are legal.
template <int ROWS, bool VEC>
__device__ __forceinline__ void load_f64(const double* col, int64_t row, double (&v)[ROWS]) {
  if constexpr (VEC && ROWS % 4 == 0) {  // 256-bit loads (PTX ISA 8.8, sm_100+); the row offset keeps them 32-byte aligned
#pragma unroll
    for (int r = 0; r < ROWS; r += 4)
      asm("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
          : "=d"(v[r]), "=d"(v[r + 1]), "=d"(v[r + 2]), "=d"(v[r + 3]) : "l"(col + row + r));
  } else if constexpr (VEC) {
#pragma unroll
    for (int r = 0; r < ROWS; r += 2) {
      const double2 t = ldg_stream_d2(col + row + r);
      v[r] = t.x;
      v[r + 1] = t.y;
    }
  } else {
#pragma unroll
    for (int r = 0; r < ROWS; ++r) v[r] = __ldg(col + row + r);
  }
}
template <int ROWS, bool VEC>
__device__ __forceinline__ void load_i32(const int* col, int64_t row, int (&v)[ROWS]) {
  if constexpr (VEC && ROWS == 8) {
    asm("ld.global.nc.L1::no_allocate.v8.s32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "l"(col + row));
  } else if constexpr (VEC && ROWS == 4) {
    const int4 t = ldg_stream_i4(col + row);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  } else if constexpr (VEC && ROWS == 2) {
    const int2 t = ldg_stream_i2(col + row);
    v[0] = t.x; v[1] = t.y;
  } else {
#pragma unroll
    for (int r = 0; r < ROWS; ++r) v[r] = __ldg(col + row + r);
  }
}

// kMin / kMax skip NaNs, which is what both of their Spark counterparts do for everything but one case:
// MultivariateOnlineSummarizer.add updates with `currMax < value` / `currMin > value` (false on NaN), and the SQL
// min()/max() behind LinearRegressionSummary.devianceResiduals order NaN above every number, so min() never returns it
// (unless nothing else is there) while max() always does -- that one case is kMaxNan.
enum ReduceKind { kSum = 0, kMin = 1, kMax = 2, kMaxNan = 3 };

__host__ __device__ __forceinline__ double combine(int kind, double a, double b) {
  if (kind == kSum) return a + b;
  if (kind == kMin) return (b != b) ? a : ((a != a) ? b : (a < b ? a : b));
  if (kind == kMax) return (b != b) ? a : ((a != a) ? b : (a > b ? a : b));
  return (a != a || b != b) ? (a + b) : (a > b ? a : b);  // kMaxNan: a NaN wins
}

__host__ __device__ constexpr double reduce_identity(int kind) {
  return kind == kMin ? __builtin_huge_val() : ((kind == kMax || kind == kMaxNan) ? -__builtin_huge_val() : 0.0);
}

__device__ __forceinline__ double warp_reduce(double v, int kind) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = combine(kind, v, __shfl_down_sync(0xffffffffu, v, off));
  return v;
}

// Deterministic block reduce of one value; result valid in thread 0.  `smem` needs kBlock/32 doubles.
__device__ __forceinline__ double block_reduce(double v, int kind, double* smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_reduce(v, kind);
  __syncthreads();  // smem may still be read from a previous call
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  if (warp == 0) {
    v = lane < (blockDim.x >> 5) ? smem[lane] : (kind == kSum ? 0.0 : smem[0]);
    v = warp_reduce(v, kind);
  }
  return v;
}

}  // namespace dq4ml
