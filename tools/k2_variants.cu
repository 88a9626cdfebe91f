// k2_variants.cu -- load-path experiments for the bench kernel (fused DQ + WLS moments over guest:int32, price:f64).
// Same arithmetic in every variant; what changes is how the 12 B/row are brought in.  Prints GB/s per variant.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o k2_variants k2_variants.cu && ./k2_variants [rows]
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

struct Acc {
  double n, sy, syy, sx, sxy, sxx;
};

__device__ __forceinline__ bool keep_row(int g, double p) {
  const double r1 = p < 20.0 ? -1.0 : p;
  const bool k1 = (r1 != r1) || (r1 > 0.0);
  const double r2 = (g < 14 && r1 > 90.0) ? -1.0 : r1;
  return k1 && ((r2 != r2) || (r2 > 0.0));
}
__device__ __forceinline__ void add_row(Acc& a, int g, double p) {
  const bool k = keep_row(g, p);
  const double x = k ? (double)g : 0.0, y = k ? p : 0.0;
  a.n += k ? 1.0 : 0.0;
  a.sy += y;
  a.syy = fma(y, y, a.syy);
  a.sx += x;
  a.sxy = fma(x, y, a.sxy);
  a.sxx = fma(x, x, a.sxx);
}
__device__ __forceinline__ int4 ld_i4(const void* p) {
  int4 r;
  asm("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ double2 ld_d2(const void* p) {
  double2 r;
  asm("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  return r;
}
// 256-bit loads (PTX ISA 8.8, sm_100+)
__device__ __forceinline__ void ld_d4(const void* p, double (&v)[4]) {
  asm("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
}
__device__ __forceinline__ void ld_i8(const void* p, int (&v)[8]) {
  asm("ld.global.nc.L1::no_allocate.v8.s32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "l"(p));
}

__device__ void finish(Acc a, double* out) {
  double v[6] = {a.n, a.sy, a.syy, a.sx, a.sxy, a.sxx};
  for (int k = 0; k < 6; ++k) {
    for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], off);
    if ((threadIdx.x & 31) == 0) atomicAdd(&out[k], v[k]);
  }
}

// A: 4 rows per thread per tile (the shipped kernel's shape)
__global__ void __launch_bounds__(256) varA(const int* g, const double* p, long long n, double* out) {
  Acc a{};
  const long long tiles = n / 1024;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long r = t * 1024 + threadIdx.x * 4;
    const int4 gv = ld_i4(g + r);
    const double2 p0 = ld_d2(p + r), p1 = ld_d2(p + r + 2);
    add_row(a, gv.x, p0.x); add_row(a, gv.y, p0.y); add_row(a, gv.z, p1.x); add_row(a, gv.w, p1.y);
  }
  finish(a, out);
}

// B: two tiles per iteration (8 rows per thread in flight)
__global__ void __launch_bounds__(256) varB(const int* g, const double* p, long long n, double* out) {
  Acc a{};
  const long long tiles = n / 2048;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long r = t * 2048 + threadIdx.x * 4;
    const int4 g0 = ld_i4(g + r), g1 = ld_i4(g + r + 1024);
    const double2 a0 = ld_d2(p + r), a1 = ld_d2(p + r + 2), b0 = ld_d2(p + r + 1024), b1 = ld_d2(p + r + 1026);
    add_row(a, g0.x, a0.x); add_row(a, g0.y, a0.y); add_row(a, g0.z, a1.x); add_row(a, g0.w, a1.y);
    add_row(a, g1.x, b0.x); add_row(a, g1.y, b0.y); add_row(a, g1.z, b1.x); add_row(a, g1.w, b1.y);
  }
  finish(a, out);
}

// C: 256-bit loads, 8 rows per thread per tile
__global__ void __launch_bounds__(256) varC(const int* g, const double* p, long long n, double* out) {
  Acc a{};
  const long long tiles = n / 2048;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long r = t * 2048 + threadIdx.x * 8;
    int gv[8];
    double p0[4], p1[4];
    ld_i8(g + r, gv);
    ld_d4(p + r, p0);
    ld_d4(p + r + 4, p1);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row(a, gv[k], p0[k]);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row(a, gv[4 + k], p1[k]);
  }
  finish(a, out);
}

// D: software-pipelined A: next tile's loads are issued before the current tile's arithmetic
__global__ void __launch_bounds__(256) varD(const int* g, const double* p, long long n, double* out) {
  Acc a{};
  const long long tiles = n / 1024;
  long long t = blockIdx.x;
  if (t >= tiles) { finish(a, out); return; }
  long long r = t * 1024 + threadIdx.x * 4;
  int4 gv = ld_i4(g + r);
  double2 p0 = ld_d2(p + r), p1 = ld_d2(p + r + 2);
  for (t += gridDim.x; t < tiles; t += gridDim.x) {
    r = t * 1024 + threadIdx.x * 4;
    const int4 gn = ld_i4(g + r);
    const double2 n0 = ld_d2(p + r), n1 = ld_d2(p + r + 2);
    add_row(a, gv.x, p0.x); add_row(a, gv.y, p0.y); add_row(a, gv.z, p1.x); add_row(a, gv.w, p1.y);
    gv = gn; p0 = n0; p1 = n1;
  }
  add_row(a, gv.x, p0.x); add_row(a, gv.y, p0.y); add_row(a, gv.z, p1.x); add_row(a, gv.w, p1.y);
  finish(a, out);
}

// E: pure read (no arithmetic beyond a checksum) -- what the memory system gives a read-only stream of this shape
__global__ void __launch_bounds__(256) varE(const int* g, const double* p, long long n, double* out) {
  double s = 0.0;
  int q = 0;
  const long long tiles = n / 1024;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long r = t * 1024 + threadIdx.x * 4;
    const int4 gv = ld_i4(g + r);
    const double2 p0 = ld_d2(p + r), p1 = ld_d2(p + r + 2);
    q ^= gv.x ^ gv.y ^ gv.z ^ gv.w;
    s += p0.x + p0.y + p1.x + p1.y;
  }
  if (s == 1.2345 && q == 77) out[0] = s;
}

// F: like A, but every CTA streams one contiguous range of tiles instead of grid-striding
__global__ void __launch_bounds__(256) varF(const int* g, const double* p, long long n, double* out) {
  Acc a{};
  const long long tiles = n / 1024;
  const long long per = (tiles + gridDim.x - 1) / gridDim.x;
  const long long t0 = blockIdx.x * per, t1 = t0 + per < tiles ? t0 + per : tiles;
  for (long long t = t0; t < t1; ++t) {
    const long long r = t * 1024 + threadIdx.x * 4;
    const int4 gv = ld_i4(g + r);
    const double2 p0 = ld_d2(p + r), p1 = ld_d2(p + r + 2);
    add_row(a, gv.x, p0.x); add_row(a, gv.y, p0.y); add_row(a, gv.z, p1.x); add_row(a, gv.w, p1.y);
  }
  finish(a, out);
}
// G: 256-bit loads, contiguous range per CTA
__global__ void __launch_bounds__(256) varG(const int* g, const double* p, long long n, double* out) {
  Acc a{};
  const long long tiles = n / 2048;
  const long long per = (tiles + gridDim.x - 1) / gridDim.x;
  const long long t0 = blockIdx.x * per, t1 = t0 + per < tiles ? t0 + per : tiles;
  for (long long t = t0; t < t1; ++t) {
    const long long r = t * 2048 + threadIdx.x * 8;
    int gv[8];
    double p0[4], p1[4];
    ld_i8(g + r, gv);
    ld_d4(p + r, p0);
    ld_d4(p + r + 4, p1);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row(a, gv[k], p0[k]);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row(a, gv[4 + k], p1[k]);
  }
  finish(a, out);
}

// ---- round 2: what separates the shipped shape (C) from the pure read (E)?
// the product's reduced predicate (three fp64 compares) and the same on the bit pattern of the price (integer compares)
__device__ __forceinline__ bool keep3(int g, double p) { return !(p < 20.0) && !(p <= 0.0) && !(g < 14 && p > 90.0); }
__device__ __forceinline__ unsigned long long f64_key(unsigned long long b) {
  if ((b << 1) == 0ULL) b = 0ULL;
  return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ bool keep_bits(int g, double p) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(p);
  if ((b & 0x7fffffffffffffffULL) > 0x7ff0000000000000ULL) return true;
  const unsigned long long k = f64_key(b);
  const unsigned long long k20 = f64_key((unsigned long long)__double_as_longlong(20.0)), k90 = f64_key((unsigned long long)__double_as_longlong(90.0));
  return !(k < k20) && !(k <= 0x8000000000000000ULL) && !(g < 14 && k > k90);
}
template <int PRED>
__device__ __forceinline__ void add_row2(Acc& a, int g, double p) {
  const bool k = PRED == 0 ? keep3(g, p) : keep_bits(g, p);
  const double x = k ? (double)g : 0.0, y = k ? p : 0.0;
  a.n += k ? 1.0 : 0.0;
  a.sy += y;
  a.syy = fma(y, y, a.syy);
  a.sx += x;
  a.sxy = fma(x, y, a.sxy);
  a.sxx = fma(x, x, a.sxx);
}
// H: C with the reduced predicate (PRED 0) or the integer predicate (PRED 1)
template <int PRED>
__global__ void __launch_bounds__(256) varH(const int* g, const double* p, long long n, double* out) {
  Acc a{};
  const long long tiles = n / 2048;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long r = t * 2048 + threadIdx.x * 8;
    int gv[8];
    double p0[4], p1[4];
    ld_i8(g + r, gv);
    ld_d4(p + r, p0);
    ld_d4(p + r + 4, p1);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row2<PRED>(a, gv[k], p0[k]);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row2<PRED>(a, gv[4 + k], p1[k]);
  }
  finish(a, out);
}
// I: H, software-pipelined: the next tile's loads are issued before this tile's arithmetic
template <int PRED>
__global__ void __launch_bounds__(256) varI(const int* g, const double* p, long long n, double* out) {
  Acc a{};
  const long long tiles = n / 2048;
  long long t = blockIdx.x;
  if (t >= tiles) { finish(a, out); return; }
  int gv[8];
  double p0[4], p1[4];
  long long r = t * 2048 + threadIdx.x * 8;
  ld_i8(g + r, gv); ld_d4(p + r, p0); ld_d4(p + r + 4, p1);
  for (t += gridDim.x; t < tiles; t += gridDim.x) {
    r = t * 2048 + threadIdx.x * 8;
    int gn[8];
    double n0[4], n1[4];
    ld_i8(g + r, gn); ld_d4(p + r, n0); ld_d4(p + r + 4, n1);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row2<PRED>(a, gv[k], p0[k]);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row2<PRED>(a, gv[4 + k], p1[k]);
#pragma unroll
    for (int k = 0; k < 8; ++k) gv[k] = gn[k];
#pragma unroll
    for (int k = 0; k < 4; ++k) { p0[k] = n0[k]; p1[k] = n1[k]; }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) add_row2<PRED>(a, gv[k], p0[k]);
#pragma unroll
  for (int k = 0; k < 4; ++k) add_row2<PRED>(a, gv[4 + k], p1[k]);
  finish(a, out);
}
// J: 16 rows per thread per tile
template <int PRED>
__global__ void __launch_bounds__(256) varJ(const int* g, const double* p, long long n, double* out) {
  Acc a{};
  const long long tiles = n / 4096;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long r = t * 4096 + threadIdx.x * 16;
    int g0[8], g1[8];
    double p0[4], p1[4], p2[4], p3[4];
    ld_i8(g + r, g0); ld_i8(g + r + 8, g1);
    ld_d4(p + r, p0); ld_d4(p + r + 4, p1); ld_d4(p + r + 8, p2); ld_d4(p + r + 12, p3);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row2<PRED>(a, g0[k], p0[k]);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row2<PRED>(a, g0[4 + k], p1[k]);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row2<PRED>(a, g1[k], p2[k]);
#pragma unroll
    for (int k = 0; k < 4; ++k) add_row2<PRED>(a, g1[4 + k], p3[k]);
  }
  finish(a, out);
}
// K: pipelined, NV 8-row vectors per thread and tile (NV = 2: 16 rows), BLK threads per CTA
template <int NV, int BLK>
__global__ void __launch_bounds__(BLK) varK(const int* g, const double* p, long long n, double* out) {
  Acc a{};
  constexpr long long kTile = static_cast<long long>(BLK) * 8 * NV;
  const long long tiles = n / kTile;
  long long t = blockIdx.x;
  if (t >= tiles) { finish(a, out); return; }
  int gv[NV][8];
  double pv[NV][8];
  auto load = [&](long long tt, int (&G)[NV][8], double (&P)[NV][8]) {
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const long long r = tt * kTile + (static_cast<long long>(v) * BLK + threadIdx.x) * 8;
      ld_i8(g + r, G[v]);
      double lo[4], hi[4];
      ld_d4(p + r, lo);
      ld_d4(p + r + 4, hi);
#pragma unroll
      for (int k = 0; k < 4; ++k) { P[v][k] = lo[k]; P[v][4 + k] = hi[k]; }
    }
  };
  auto compute = [&](const int (&G)[NV][8], const double (&P)[NV][8]) {
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
      for (int k = 0; k < 8; ++k) add_row2<0>(a, G[v][k], P[v][k]);
  };
  load(t, gv, pv);
  for (t += gridDim.x; t < tiles; t += gridDim.x) {
    int gn[NV][8];
    double pn[NV][8];
    load(t, gn, pn);
    compute(gv, pv);
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
      for (int k = 0; k < 8; ++k) { gv[v][k] = gn[v][k]; pv[v][k] = pn[v][k]; }
  }
  compute(gv, pv);
  finish(a, out);
}
template <class K>
void run_blk(const char* name, K kernel, int per_sm, int blk, const int* g, const double* p, long long n, double* out);

// E8: pure read with the 256-bit loads of C
__global__ void __launch_bounds__(256) varE8(const int* g, const double* p, long long n, double* out) {
  double s = 0.0;
  int q = 0;
  const long long tiles = n / 2048;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long r = t * 2048 + threadIdx.x * 8;
    int gv[8];
    double p0[4], p1[4];
    ld_i8(g + r, gv); ld_d4(p + r, p0); ld_d4(p + r + 4, p1);
#pragma unroll
    for (int k = 0; k < 8; ++k) q ^= gv[k];
#pragma unroll
    for (int k = 0; k < 4; ++k) s += p0[k] + p1[k];
  }
  if (s == 1.2345 && q == 77) out[0] = s;
}

template <class K>
void run(const char* name, K kernel, int per_sm, const int* g, const double* p, long long n, double* out) {
  run_blk(name, kernel, per_sm, 256, g, p, n, out);
}
template <class K>
void run_blk(const char* name, K kernel, int per_sm, int blk, const int* g, const double* p, long long n, double* out) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f, sum = 0.f;
  const int reps = 20;
  for (int i = 0; i < 3; ++i) kernel<<<sms * per_sm, blk>>>(g, p, n, out);
  for (int i = 0; i < reps; ++i) {
    cudaMemsetAsync(out, 0, 64);
    cudaEventRecord(e0);
    kernel<<<sms * per_sm, blk>>>(g, p, n, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    best = ms < best ? ms : best;
    sum += ms;
  }
  double h[6];
  cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
  printf("{\"variant\": \"%s\", \"ctas_per_sm\": %d, \"best_ms\": %.4f, \"avg_ms\": %.4f, \"best_gbs\": %.1f, \"avg_gbs\": %.1f, \"count\": %.0f, \"err\": \"%s\"}\n",
         name, per_sm, best, sum / reps, 12.0 * n / best / 1e6, 12.0 * n / (sum / reps) / 1e6, h[0], cudaGetErrorString(cudaGetLastError()));
}

__global__ void fill(int* g, double* p, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    unsigned long long z = (i + 1) * 0x9E3779B97F4A7C15ull;
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27;
    g[i] = 1 + (int)(z % 35);
    p[i] = 5.0 * g[i] + 22.0 + (double)((z >> 20) & 1023) * 0.01 - ((z >> 40) % 100 == 0 ? 200.0 : 0.0);
  }
}

int main(int argc, char** argv) {
  const long long n = argc > 1 ? atoll(argv[1]) : 100000000LL;
  int* g;
  double *p, *out;
  cudaMalloc(&g, n * 4);
  cudaMalloc(&p, n * 8);
  cudaMalloc(&out, 64);
  fill<<<148 * 8, 256>>>(g, p, n);
  cudaDeviceSynchronize();
  if (argc > 3) {  // pipelined shapes
    for (int per_sm : {1, 2, 3, 4}) {
      run_blk("K_pipe_8rows_256thr", varK<1, 256>, per_sm, 256, g, p, n, out);
      run_blk("K_pipe_16rows_256thr", varK<2, 256>, per_sm, 256, g, p, n, out);
      run_blk("K_pipe_24rows_256thr", varK<3, 256>, per_sm, 256, g, p, n, out);
      run_blk("K_pipe_8rows_512thr", varK<1, 512>, per_sm, 512, g, p, n, out);
      run_blk("K_pipe_16rows_512thr", varK<2, 512>, per_sm, 512, g, p, n, out);
      run_blk("K_pipe_32rows_256thr", varK<4, 256>, per_sm, 256, g, p, n, out);
      run_blk("K_pipe_24rows_384thr", varK<3, 384>, per_sm, 384, g, p, n, out);
      run_blk("K_pipe_32rows_128thr", varK<4, 128>, per_sm, 128, g, p, n, out);
      run_blk("K_pipe_48rows_128thr", varK<6, 128>, per_sm, 128, g, p, n, out);
      run_blk("K_pipe_24rows_192thr", varK<3, 192>, per_sm, 192, g, p, n, out);
      run_blk("K_pipe_8rows_384thr", varK<1, 384>, per_sm, 384, g, p, n, out);
      run_blk("K_pipe_16rows_384thr", varK<2, 384>, per_sm, 384, g, p, n, out);
      run("E8_pure_read_256bit", varE8, per_sm, g, p, n, out);
      run("C_256bit", varC, per_sm, g, p, n, out);
    }
    return 0;
  }
  if (argc > 2) {  // round 2 variants only
    for (int per_sm : {2, 3, 4, 5, 6}) {
      run("C_256bit", varC, per_sm, g, p, n, out);
      run("E8_pure_read_256bit", varE8, per_sm, g, p, n, out);
      run("H_pred3", varH<0>, per_sm, g, p, n, out);
      run("H_pred_int", varH<1>, per_sm, g, p, n, out);
      run("I_pipelined_pred3", varI<0>, per_sm, g, p, n, out);
      run("I_pipelined_pred_int", varI<1>, per_sm, g, p, n, out);
      run("J_16rows_pred3", varJ<0>, per_sm, g, p, n, out);
      run("J_16rows_pred_int", varJ<1>, per_sm, g, p, n, out);
    }
    return 0;
  }
  for (int per_sm : {2, 3, 4, 5, 6, 7, 8}) {
    run("A_4rows", varA, per_sm, g, p, n, out);
    run("C_256bit", varC, per_sm, g, p, n, out);
    run("E_pure_read", varE, per_sm, g, p, n, out);
    run("F_4rows_chunked", varF, per_sm, g, p, n, out);
    run("G_256bit_chunked", varG, per_sm, g, p, n, out);
  }
  return 0;
}
