// Microbenchmark (tuning aid, not product code): FP64 FMA dependent-chain latency and per-warp issue interval on this GPU,
// shared-memory load-to-use latency, for 1 warp per SM and for W warps per SMSP.
#include <cstdio>
#include <cuda_runtime.h>
template <int CH>
__global__ void chain(double* out, long long* clk, int iters, double a, double b) {
  double x[CH];
  for (int i = 0; i < CH; i++) x[i] = a + i + threadIdx.x;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < CH; i++) x[i] = fma(x[i], b, a);
  }
  long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < CH; i++) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
}
__global__ void lds_chain(double* out, long long* clk, int iters) {
  __shared__ int idx[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) idx[i] = (i + 33) & 1023;
  __syncthreads();
  int j = threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) j = idx[j];
  long long t1 = clock64();
  out[threadIdx.x] = j;
  if (threadIdx.x == 0) clk[0] = t1 - t0;
}
int main() {
  double* out; long long* clk; cudaMalloc(&out, 1 << 24); cudaMalloc(&clk, 64);
  long long h;
  const int iters = 4096;
#define RUN(CH, THREADS)                                                                     \
  chain<CH><<<1, THREADS>>>(out, clk, iters, 1.0, 0.999);                                    \
  cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);                                            \
  printf("DFMA chains/thread %2d  threads %4d : %.2f clk per chain step, %.2f clk per warp-DFMA per SMSP\n", CH, THREADS, \
         (double)h / iters, (double)h / iters / CH / ((THREADS + 127) / 128));
  RUN(1, 32) RUN(2, 32) RUN(4, 32) RUN(8, 32) RUN(16, 32) RUN(1, 128) RUN(4, 128) RUN(8, 128) RUN(8, 512) RUN(8, 1024)
  lds_chain<<<1, 32>>>(out, clk, iters);
  cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
  printf("LDS dependent chain: %.2f clk per load\n", (double)h / iters);
  return 0;
}
