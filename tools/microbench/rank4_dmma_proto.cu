// Prototype for round 2 (tuning aid, not product code): the measurement update's covariance step  P += U V^T  (inner dimension
// 2r = 4 for the FEAT / ACC-drag models: U = [W, -K], V = [K, H], vi_ekf_meas.cpp:254-265 in rank form) with P held as
// mma.m8n8k4 f64 ACCUMULATOR FRAGMENTS in registers: n = 96 (N = 25 padded: 91 -> 96) = 78 lower-triangle 8x8 blocks spread over
// the 3 warps of a filter (26 blocks = 52 accumulator doubles per lane), U and V (n x 4 each) in shared memory.  One DMMA per
// block per update.  Reports updates/s and checks the result against the same sum evaluated with plain FMAs on the host.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
constexpr int NB = 12, NBLK = NB * (NB + 1) / 2, W = 3, PER = NBLK / W, N = 8 * NB;
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int FPC>
__global__ void __launch_bounds__(32 * W * FPC) rank4_kernel(double* P, const double* UV, int filters, int updates) {
  __shared__ double sU[FPC][N * 4], sV[FPC][N * 4];
  __shared__ int sOff[NBLK][2];
  const int g = threadIdx.x / (32 * W), t = threadIdx.x % (32 * W), w = t / 32, lane = t % 32;
  for (int blk = threadIdx.x; blk < NBLK; blk += blockDim.x) {
    int bi = 0;
    while ((bi + 1) * (bi + 2) / 2 <= blk) bi++;
    sOff[blk][0] = 8 * bi * 4;
    sOff[blk][1] = 8 * (blk - bi * (bi + 1) / 2) * 4;
  }
  __syncthreads();
  for (int f = blockIdx.x * FPC + g; f < filters; f += gridDim.x * FPC) {
    double c0[PER], c1[PER];
    double* Pf = P + (size_t)f * NBLK * 64;
#pragma unroll
    for (int k = 0; k < PER; k++) {  // fragment-major storage: block (w + W k), lane -> 2 doubles
      const double2 v = *(const double2*)(Pf + ((w + W * k) * 32 + lane) * 2);
      c0[k] = v.x;
      c1[k] = v.y;
    }
    for (int i = t; i < N * 4; i += 32 * W) {
      sU[g][i] = UV[i];
      sV[g][i] = UV[N * 4 + i];
    }
    asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(32 * W) : "memory");
    const int lo = (lane >> 2) * 4 + (lane & 3);
    for (int u = 0; u < updates; u++) {
#pragma unroll
      for (int k = 0; k < PER; k++) {
        const int blk = w + W * k;
        const double a = sU[g][sOff[blk][0] + lo];
        const double b = sV[g][sOff[blk][1] + lo];
        dmma(c0[k], c1[k], a, b);
      }
    }
#pragma unroll
    for (int k = 0; k < PER; k++) *(double2*)(Pf + ((w + W * k) * 32 + lane) * 2) = make_double2(c0[k], c1[k]);
    asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(32 * W) : "memory");
  }
}
int main() {
  const int filters = 148 * 4 * 64, updates = 200;
  cudaDeviceProp pr;
  cudaGetDeviceProperties(&pr, 0);
  std::vector<double> hUV(N * 8), hP((size_t)NBLK * 64, 0.0);
  srand(1);
  for (auto& v : hUV) v = (rand() / (double)RAND_MAX - 0.5) * 0.1;
  double *dP, *dUV;
  cudaMalloc(&dP, sizeof(double) * NBLK * 64 * (size_t)filters);
  cudaMalloc(&dUV, sizeof(double) * N * 8);
  cudaMemcpy(dUV, hUV.data(), sizeof(double) * N * 8, cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float ms = 0;
  constexpr int FPC = 4;
  for (int rep = 0; rep < 3; rep++) {
    cudaMemset(dP, 0, sizeof(double) * NBLK * 64 * (size_t)filters);
    cudaEventRecord(e0);
    rank4_kernel<FPC><<<pr.multiProcessorCount * 2, 32 * W * FPC>>>(dP, dUV, filters, updates);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
  }
  printf("launch: %s\n", cudaGetErrorString(cudaGetLastError()));
  cudaMemcpy(hP.data(), dP, sizeof(double) * NBLK * 64, cudaMemcpyDeviceToHost);
  double worst = 0;
  for (int blk = 0; blk < NBLK; blk++) {
    int bi = 0;
    while ((bi + 1) * (bi + 2) / 2 <= blk) bi++;
    const int bj = blk - bi * (bi + 1) / 2;
    for (int lane = 0; lane < 32; lane++)
      for (int e = 0; e < 2; e++) {
        const int r = 8 * bi + (lane >> 2), c = 8 * bj + 2 * (lane & 3) + e;
        double s = 0;
        for (int k = 0; k < 4; k++) s = fma(hUV[r * 4 + k], hUV[N * 4 + c * 4 + k], s);
        const double ref = s * updates, got = hP[(blk * 32 + lane) * 2 + e];
        worst = fmax(worst, fabs(got - ref) / fmax(fabs(ref), 1e-30));
      }
  }
  const double ups = (double)filters * updates / (ms * 1e-3);
  printf("rank-4 update of a 96x96 symmetric P on DMMA fragments: %.1f M updates/s on %d SMs (%.2f ms, %d filters x %d updates), "
         "%.2f TFLOP/s of DMMA work, worst rel err vs FMA sum %.2e\n",
         ups * 1e-6, pr.multiProcessorCount, ms, filters, updates, ups * NBLK * 512.0 * 1e-12, worst);
  printf("for scale: the round-1 tile executor applies ~34 M updates/s inside 6.97 M filter-steps/s (4.9 updates per step)\n");
  return 0;
}
