// filter_ceiling.cu -- what the access pattern of the DQ filter kernel can reach without its logic: copies of
// guest:int32[n] + price:f64[n] (12 B read + 12 B written per row, like the filter at ~100 % survival) with
//   A: the filter's own pattern -- 4096-row tiles, 256 threads, lane l owns rows l, l+32, ... (scalar 4 B / 8 B accesses),
//      straight from registers to global memory;
//   B: A, but parked in shared memory first and written out tile-wide (the filter's write-out loop);
//   C: 8 consecutive rows per thread with 256-bit loads and stores (what a compaction-free copy would do);
//   D: A's loads, C-style 128-bit stores out of shared memory.
// Prints GB/s (read + write) per variant.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o filter_ceiling filter_ceiling.cu && ./filter_ceiling [rows]
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

constexpr int kT = 256, kRows = 16;
constexpr long long kTile = kT * kRows;

__global__ void __launch_bounds__(kT, 4) copyA(const int* __restrict__ g, const double* __restrict__ p, int* go, double* po) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long r0 = blockIdx.x * kTile + warp * (32 * kRows) + lane;
  int gv[kRows];
  double pv[kRows];
#pragma unroll
  for (int j = 0; j < kRows; ++j) {
    gv[j] = __ldg(g + r0 + 32 * j);
    pv[j] = __ldg(p + r0 + 32 * j);
  }
#pragma unroll
  for (int j = 0; j < kRows; ++j) {
    go[r0 + 32 * j] = gv[j];
    po[r0 + 32 * j] = pv[j];
  }
}

__global__ void __launch_bounds__(kT, 4) copyB(const int* __restrict__ g, const double* __restrict__ p, int* go, double* po) {
  extern __shared__ __align__(16) unsigned char sm[];
  double* sp = reinterpret_cast<double*>(sm);
  int* sg = reinterpret_cast<int*>(sm + kTile * sizeof(double));
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long t0 = blockIdx.x * kTile;
  const int l0 = warp * (32 * kRows) + lane;
#pragma unroll
  for (int j = 0; j < kRows; ++j) {
    sg[l0 + 32 * j] = __ldg(g + t0 + l0 + 32 * j);
    sp[l0 + 32 * j] = __ldg(p + t0 + l0 + 32 * j);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kTile; i += kT) {
    go[t0 + i] = sg[i];
    po[t0 + i] = sp[i];
  }
}

__global__ void __launch_bounds__(kT, 4) copyC(const int* __restrict__ g, const double* __restrict__ p, int* go, double* po) {
  const long long r = (static_cast<long long>(blockIdx.x) * kT + threadIdx.x) * 8;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const long long rr = r + static_cast<long long>(k) * gridDim.x * kT * 8;
    int gv[8];
    double pv[8];
    asm("ld.global.nc.L1::no_allocate.v8.s32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(gv[0]), "=r"(gv[1]), "=r"(gv[2]), "=r"(gv[3]), "=r"(gv[4]), "=r"(gv[5]), "=r"(gv[6]), "=r"(gv[7]) : "l"(g + rr));
    asm("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(pv[0]), "=d"(pv[1]), "=d"(pv[2]), "=d"(pv[3]) : "l"(p + rr));
    asm("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(pv[4]), "=d"(pv[5]), "=d"(pv[6]), "=d"(pv[7]) : "l"(p + rr + 4));
    asm volatile("st.global.v8.s32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(go + rr), "r"(gv[0]), "r"(gv[1]), "r"(gv[2]), "r"(gv[3]),
                 "r"(gv[4]), "r"(gv[5]), "r"(gv[6]), "r"(gv[7]) : "memory");
    asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(po + rr), "d"(pv[0]), "d"(pv[1]), "d"(pv[2]), "d"(pv[3]) : "memory");
    asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(po + rr + 4), "d"(pv[4]), "d"(pv[5]), "d"(pv[6]), "d"(pv[7]) : "memory");
  }
}

__global__ void __launch_bounds__(kT, 4) copyD(const int* __restrict__ g, const double* __restrict__ p, int* go, double* po) {
  extern __shared__ __align__(16) unsigned char sm[];
  double* sp = reinterpret_cast<double*>(sm);
  int* sg = reinterpret_cast<int*>(sm + kTile * sizeof(double));
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long t0 = blockIdx.x * kTile;
  const int l0 = warp * (32 * kRows) + lane;
#pragma unroll
  for (int j = 0; j < kRows; ++j) {
    sg[l0 + 32 * j] = __ldg(g + t0 + l0 + 32 * j);
    sp[l0 + 32 * j] = __ldg(p + t0 + l0 + 32 * j);
  }
  __syncthreads();
  for (int i = threadIdx.x * 4; i < kTile; i += kT * 4) *reinterpret_cast<int4*>(go + t0 + i) = *reinterpret_cast<const int4*>(sg + i);
  for (int i = threadIdx.x * 2; i < kTile; i += kT * 2) *reinterpret_cast<double2*>(po + t0 + i) = *reinterpret_cast<const double2*>(sp + i);
}

int main(int argc, char** argv) {
  long long n = argc > 1 ? atoll(argv[1]) : 1000000000LL;
  n = n / (2 * kTile) * (2 * kTile);
  int *g, *go;
  double *p, *po;
  cudaMalloc(&g, n * 4);
  cudaMalloc(&go, n * 4);
  cudaMalloc(&p, n * 8);
  cudaMalloc(&po, n * 8);
  cudaMemset(g, 1, n * 4);
  cudaMemset(p, 1, n * 8);
  const size_t smem = kTile * 12;
  cudaFuncSetAttribute(copyB, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  cudaFuncSetAttribute(copyD, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const unsigned tiles = static_cast<unsigned>(n / kTile);
  for (int v = 0; v < 4; ++v) {
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
      cudaEventRecord(e0);
      if (v == 0) copyA<<<tiles, kT>>>(g, p, go, po);
      if (v == 1) copyB<<<tiles, kT, smem>>>(g, p, go, po);
      if (v == 2) copyC<<<static_cast<unsigned>(n / (kT * 16)), kT>>>(g, p, go, po);
      if (v == 3) copyD<<<tiles, kT, smem>>>(g, p, go, po);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep > 0 && ms < best) best = ms;
    }
    printf("{\"variant\": \"%c\", \"rows\": %lld, \"ms\": %.4f, \"gbs\": %.1f, \"err\": \"%s\"}\n", "ABCD"[v], n, best, 24.0 * n / best / 1e6,
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
