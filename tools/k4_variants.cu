// k4_variants.cu -- how far is the 16-column loss/gradient pass (config 3: 1 int32 + 15 fp64 feature columns + the
// label = 132 B/row) from what the memory system gives a read-only stream of that shape?
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o k4_variants k4_variants.cu && ./k4_variants [rows]
// Variants: R = pure read (checksum only) with 4 rows x 16 columns per thread (the product's register shape), at
// several occupancies; L = the same loads with the LeastSquaresAggregator arithmetic (dot product, loss, 16 gradient
// sums), one CTA per SM; Q = columns dealt out to 4 lanes (4 columns x 8 rows per lane), software-pipelined: the next
// tile's loads are issued before this tile's arithmetic.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

constexpr int D = 16;
struct Cols {
  const void* p[D];
  const double* y;
};

__device__ __forceinline__ void ld_d4(const double* p, double (&v)[4]) {
  asm("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
}
__device__ __forceinline__ int4 ld_i4(const int* p) {
  int4 r;
  asm("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void ld_i8(const int* p, int (&v)[8]) {
  asm("ld.global.nc.L1::no_allocate.v8.s32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "l"(p));
}

// R: pure read, 4 rows x 16 columns + label per thread and tile
template <int BLK>
__global__ void __launch_bounds__(BLK) varR(const Cols c, long long n, double* out) {
  double s = 0.0;
  const long long tile_rows = BLK * 4, tiles = n / tile_rows;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long r = t * tile_rows + threadIdx.x * 4;
    const int4 g = ld_i4(static_cast<const int*>(c.p[0]) + r);
    s += g.x + g.y + g.z + g.w;
#pragma unroll
    for (int j = 1; j < D; ++j) {
      double v[4];
      ld_d4(static_cast<const double*>(c.p[j]) + r, v);
      s += v[0] + v[1] + v[2] + v[3];
    }
    double v[4];
    ld_d4(c.y + r, v);
    s += v[0] + v[1] + v[2] + v[3];
  }
  if (s == 1.2345) out[0] = s;
}

// L: the product's shape with the arithmetic of LsqOp<16>
__global__ void __launch_bounds__(256, 1) varL(const Cols c, long long n, double* out) {
  double acc[2 + D];
#pragma unroll
  for (int k = 0; k < 2 + D; ++k) acc[k] = 0.0;
  const long long tile_rows = 256 * 4, tiles = n / tile_rows;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long r = t * tile_rows + threadIdx.x * 4;
    double x[D][4], y[4];
    const int4 g = ld_i4(static_cast<const int*>(c.p[0]) + r);
    x[0][0] = g.x; x[0][1] = g.y; x[0][2] = g.z; x[0][3] = g.w;
#pragma unroll
    for (int j = 1; j < D; ++j) ld_d4(static_cast<const double*>(c.p[j]) + r, x[j]);
    ld_d4(c.y + r, y);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const bool keep = !(y[q] < 20.0) && !(y[q] <= 0.0) && !(static_cast<int>(x[0][q]) < 14 && y[q] > 90.0);
      double dot = 0.0;
#pragma unroll
      for (int j = 0; j < D; ++j) dot = fma(keep ? x[j][q] : 0.0, 0.01 * (j + 1), dot);
      const double diff = keep ? dot - y[q] * 0.02 + 0.3 : 0.0;
      acc[0] = fma(0.5 * diff, diff, acc[0]);
      acc[1] += keep ? 1.0 : 0.0;
#pragma unroll
      for (int j = 0; j < D; ++j) acc[2 + j] = fma(diff, keep ? x[j][q] : 0.0, acc[2 + j]);
    }
  }
  for (int k = 0; k < 2 + D; ++k) {
    double v = acc[k];
    for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
    if ((threadIdx.x & 31) == 0) atomicAdd(&out[k], v);
  }
}

// Q: 4 lanes per row group, lane h holds columns 4 h .. 4 h + 3 of PR rows; pipelined (PIPE) or not
template <int PR, int BLK, bool PIPE>
__global__ void __launch_bounds__(BLK, 1) varQ(const Cols c, long long n, double* out) {
  const int lane = threadIdx.x & 31, h = lane & 3;
  double acc[2 + 4];
#pragma unroll
  for (int k = 0; k < 6; ++k) acc[k] = 0.0;
  double cst[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) cst[j] = 0.01 * (4 * h + j + 1);
  constexpr long long kTile = static_cast<long long>(BLK / 4) * PR;
  const long long tiles = n / kTile;
  const long long in_tile = static_cast<long long>(threadIdx.x / 4) * PR;
  auto load = [&](long long t, double (&x)[4][PR], double (&y)[PR]) {
    const long long r = t * kTile + in_tile;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int col = 4 * h + j;
      if (col == 0) {
        int g[PR];
        if (PR == 8) ld_i8(static_cast<const int*>(c.p[0]) + r, reinterpret_cast<int(&)[8]>(g));
        else { const int4 q = ld_i4(static_cast<const int*>(c.p[0]) + r); g[0] = q.x; g[1] = q.y; g[2] = q.z; g[3] = q.w; }
#pragma unroll
        for (int q = 0; q < PR; ++q) x[j][q] = g[q];
      } else {
#pragma unroll
        for (int q = 0; q < PR; q += 4) ld_d4(static_cast<const double*>(c.p[col]) + r + q, reinterpret_cast<double(&)[4]>(x[j][q]));
      }
    }
#pragma unroll
    for (int q = 0; q < PR; q += 4) ld_d4(c.y + r + q, reinterpret_cast<double(&)[4]>(y[q]));
  };
  auto compute = [&](const double (&x)[4][PR], const double (&y)[PR]) {
#pragma unroll
    for (int q = 0; q < PR; ++q) {
      const double g = __shfl_sync(0xffffffffu, x[0][q], lane & ~3);
      const bool keep = !(y[q] < 20.0) && !(y[q] <= 0.0) && !(static_cast<int>(g) < 14 && y[q] > 90.0);
      double part = 0.0;
#pragma unroll
      for (int j = 0; j < 4; ++j) part = fma(keep ? x[j][q] : 0.0, cst[j], part);
      part += __shfl_xor_sync(0xffffffffu, part, 1);
      part += __shfl_xor_sync(0xffffffffu, part, 2);
      const double diff = keep ? part - y[q] * 0.02 + 0.3 : 0.0;
      acc[0] = fma(0.5 * diff, diff, acc[0]);
      acc[1] += keep ? 1.0 : 0.0;
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[2 + j] = fma(diff, keep ? x[j][q] : 0.0, acc[2 + j]);
    }
  };
  long long t = blockIdx.x;
  if (t < tiles) {
    double x[4][PR], y[PR];
    load(t, x, y);
    if (PIPE) {
      for (long long nx = t + gridDim.x; nx < tiles; nx += gridDim.x) {
        double xn[4][PR], yn[PR];
        load(nx, xn, yn);
        compute(x, y);
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int q = 0; q < PR; ++q) x[j][q] = xn[j][q];
#pragma unroll
        for (int q = 0; q < PR; ++q) y[q] = yn[q];
      }
      compute(x, y);
    } else {
      compute(x, y);
      for (long long nx = t + gridDim.x; nx < tiles; nx += gridDim.x) {
        load(nx, x, y);
        compute(x, y);
      }
    }
  }
  for (int k = 0; k < 6; ++k) {
    double v = acc[k];
    for (int off = 16; off >= 4; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
    if (lane < 4) atomicAdd(&out[k < 2 ? k : 2 + 4 * lane + (k - 2)], (k < 2 && lane != 0) ? 0.0 : v);
  }
}

template <class K>
void run(const char* name, K kernel, int per_sm, int blk, const Cols& c, long long n, double* out) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f, sum = 0.f;
  const int reps = 10;
  for (int i = 0; i < 2; ++i) kernel<<<sms * per_sm, blk>>>(c, n, out);
  for (int i = 0; i < reps; ++i) {
    cudaMemsetAsync(out, 0, 256);
    cudaEventRecord(e0);
    kernel<<<sms * per_sm, blk>>>(c, n, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    best = ms < best ? ms : best;
    sum += ms;
  }
  double hcount[2];
  cudaMemcpy(hcount, out, sizeof(hcount), cudaMemcpyDeviceToHost);
  const double bytes = (4.0 + 8.0 * (D - 1) + 8.0) * n;
  printf("{\"variant\": \"%s\", \"ctas_per_sm\": %d, \"threads\": %d, \"best_ms\": %.4f, \"avg_ms\": %.4f, \"best_gbs\": %.1f, \"avg_gbs\": %.1f, \"count\": %.0f, \"err\": \"%s\"}\n",
         name, per_sm, blk, best, sum / reps, bytes / best / 1e6, bytes / (sum / reps) / 1e6, hcount[1], cudaGetErrorString(cudaGetLastError()));
}

__global__ void fill(int* g, double* p, double* x, long long n, long long stride) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    unsigned long long z = (i + 1) * 0x9E3779B97F4A7C15ull;
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27;
    g[i] = 1 + (int)(z % 35);
    p[i] = 5.0 * g[i] + 22.0 + (double)((z >> 20) & 1023) * 0.01 - ((z >> 40) % 100 == 0 ? 200.0 : 0.0);
    for (int j = 0; j < D - 1; ++j) x[j * stride + i] = (double)((z >> (j + 3)) & 255) * 0.01;
  }
}

int main(int argc, char** argv) {
  const long long n = argc > 1 ? atoll(argv[1]) : 200000000LL;
  int* g;
  double *p, *x, *out;
  cudaMalloc(&g, n * 4);
  cudaMalloc(&p, n * 8);
  cudaMalloc(&x, n * 8 * (D - 1));
  cudaMalloc(&out, 256);
  fill<<<148 * 8, 256>>>(g, p, x, n, n);
  cudaDeviceSynchronize();
  Cols c;
  c.p[0] = g;
  for (int j = 1; j < D; ++j) c.p[j] = x + (long long)(j - 1) * n;
  c.y = p;
  for (int per_sm : {1, 2, 3}) {
    run("R_pure_read_256thr", varR<256>, per_sm, 256, c, n, out);
    run("R_pure_read_512thr", varR<512>, per_sm, 512, c, n, out);
  }
  run("L_lsq_4x16_256thr", varL, 1, 256, c, n, out);
  run("Q_quad_4rows_512thr", varQ<4, 512, false>, 1, 512, c, n, out);
  run("Q_quad_4rows_512thr_pipe", varQ<4, 512, true>, 1, 512, c, n, out);
  run("Q_quad_8rows_512thr", varQ<8, 512, false>, 1, 512, c, n, out);
  run("Q_quad_8rows_384thr_pipe", varQ<8, 384, true>, 1, 384, c, n, out);
  run("Q_quad_8rows_256thr_pipe", varQ<8, 256, true>, 1, 256, c, n, out);
  run("Q_quad_8rows_320thr_pipe", varQ<8, 320, true>, 1, 320, c, n, out);
  run("Q_quad_4rows_640thr_pipe", varQ<4, 640, true>, 1, 640, c, n, out);
  run("Q_quad_4rows_768thr_pipe", varQ<4, 768, true>, 1, 768, c, n, out);
  return 0;
}
