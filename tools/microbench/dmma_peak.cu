// Microbenchmark (tuning aid, not product code): FP64 tensor-core MMA (mma.sync m8n8k4 f64, SASS DMMA) throughput and
// dependent-chain latency on this GPU next to the DFMA pipe, to decide whether the O(N^2) covariance products should
// move to DMMA fragments (north_star: "only where ncu shows it beats CUDA-core FP64").
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int CH>
__global__ void dmma_chain(double* out, long long* clk, int iters, double a, double b) {
  double c0[CH], c1[CH];
  for (int i = 0; i < CH; i++) { c0[i] = threadIdx.x + i; c1[i] = i; }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < CH; i++) dmma(c0[i], c1[i], a, b);
  }
  long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < CH; i++) s += c0[i] + c1[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
}
template <int CH>
__global__ void dfma_chain(double* out, int iters, double a, double b) {
  double x[CH];
  for (int i = 0; i < CH; i++) x[i] = a + i + threadIdx.x;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < CH; i++) x[i] = fma(x[i], b, a);
  }
  double s = 0;
  for (int i = 0; i < CH; i++) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  double* out; long long* clk; cudaMalloc(&out, 1 << 26); cudaMalloc(&clk, 64);
  cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
  long long h; const int iters = 4096;
#define LAT(CH, THREADS)                                                                              \
  dmma_chain<CH><<<1, THREADS>>>(out, clk, iters, 1e-3, 0.5);                                         \
  cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);                                                     \
  printf("DMMA chains/warp %2d  threads %4d : %.2f clk per chain step, %.2f clk per DMMA per SMSP (256 FMA each)\n", CH, THREADS, \
         (double)h / iters, (double)h / iters / CH / ((THREADS + 127) / 128));
  LAT(1, 32) LAT(2, 32) LAT(4, 32) LAT(8, 32) LAT(16, 32) LAT(8, 128) LAT(8, 256) LAT(8, 512) LAT(16, 1024)
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1); float ms;
  const int big = 1 << 15;
  for (int rep = 0; rep < 2; rep++) {
    cudaEventRecord(e0);
    dmma_chain<16><<<pr.multiProcessorCount * 2, 512>>>(out, clk, big, 1e-3, 0.5);
    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
  }
  double fl = 2.0 * 256 * 16 * (double)big * (512 / 32) * pr.multiProcessorCount * 2;
  printf("DMMA m8n8k4 device throughput: %.2f TFLOP/s (%d SMs, %.1f ms)\n", fl / ms * 1e-9, pr.multiProcessorCount, ms);
  for (int rep = 0; rep < 2; rep++) {
    cudaEventRecord(e0);
    dfma_chain<16><<<pr.multiProcessorCount * 2, 512>>>(out, big, 1.0, 0.999);
    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
  }
  fl = 2.0 * 16 * (double)big * 512 * pr.multiProcessorCount * 2;
  printf("DFMA device throughput: %.2f TFLOP/s (%.1f ms)\n", fl / ms * 1e-9, ms);
  return 0;
}
