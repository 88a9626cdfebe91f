// fp64_peak.cu -- microbenchmark of the two fp64 pipes a Gram kernel could use on B200 (sm_100a):
// DFMA (CUDA cores) and DMMA (mma.sync.m8n8k4.f64).  Prints achieved TFLOP/s; the larger one is the roofline
// denominator quoted for the d = 256 normal-equation pass (BASELINE.md asks the builder to measure it).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_peak fp64_peak.cu && ./fp64_peak
#include <cuda_runtime.h>
#include <cstdio>

__global__ void __launch_bounds__(256) dfma_kernel(double* out, int iters) {
  double a[16], x = 1.0 + threadIdx.x * 1e-9, y = 0.999999;
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = fma(a[i], x, y);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(256) dmma_kernel(double* out, int iters) {
  double c[16][2];
#pragma unroll
  for (int i = 0; i < 16; ++i) c[i][0] = c[i][1] = 0.0;
  double a = 1.0 + threadIdx.x * 1e-9, b = 0.5;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) dmma884(c[i][0], c[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

#ifdef TRY_M16
__device__ __forceinline__ void dmma16816(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}
__global__ void __launch_bounds__(256) dmma16_kernel(double* out, int iters) {
  double c[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.0;
  double a[8], b[4];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = 1.0 + threadIdx.x * 1e-9 + i;
#pragma unroll
  for (int i = 0; i < 4; ++i) b[i] = 0.5 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dmma16816(c[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
#endif

template <class K>
double run(K kernel, double flops_per_thread_iter, int blocks_per_sm, const char* name) {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = sms * blocks_per_sm, iters = 20000;
  double* out;
  cudaMalloc(&out, sizeof(double) * grid * 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  kernel<<<grid, 256>>>(out, 1000);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    kernel<<<grid, 256>>>(out, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  const double tf = flops_per_thread_iter * iters * grid * 256.0 / (best * 1e-3) / 1e12;
  printf("{\"pipe\": \"%s\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f, \"err\": \"%s\"}\n", name, blocks_per_sm, best, tf,
         cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
  return tf;
}

int main() {
  for (int b : {1, 2, 4, 8}) run(dfma_kernel, 16 * 2.0, b, "DFMA");
  // one warp-level m8n8k4 = 8*8*4*2 = 512 flop -> 16 flop per thread
  for (int b : {1, 2, 4, 8}) run(dmma_kernel, 16 * 16.0, b, "DMMA.m8n8k4");
#ifdef TRY_M16
  // m16n8k16 = 16*8*16*2 = 4096 flop per warp -> 128 per thread
  for (int b : {1, 2, 4, 8}) run(dmma16_kernel, 8 * 128.0, b, "DMMA.m16n8k16");
#endif
  return 0;
}
