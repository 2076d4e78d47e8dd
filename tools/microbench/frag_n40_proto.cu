// Prototype (tuning aid, not product code) for the executor planned in DESIGN.md 7: ONE WARP PER FILTER with the whole n = 40
// covariance (N = 8: 16 + 3 x 8) as 5 x 5 blocks of mma.sync.m8n8k4.f64 ACCUMULATOR FRAGMENTS in registers (50 doubles per lane, both
// triangles), no cross-warp synchronisation.  One "step" = one covariance propagate  P' = A P A^T  (A dense 40 x 40 from shared
// memory: the worst case, the real A = I + M has 17 of its 25 blocks non-zero) + UPD rank-4 updates  P += G K^T + K G^T  (the
// symmetric Joseph form of viekf_octet.cu), G, K (40 x 2 each) from shared memory.
//   T = A P    : block (I,J) = sum_K A(I,K) P(K,J); the B operand P(K,J)[4kk+t][g] is read from the mirror block (J,K), element
//                (g, 4kk+t), which lives in lane 4g + 2kk + t/2, register t%2: an intra-quad shuffle pair + select
//   P' = T A^T : block (I,J) = sum_K T(I,K) A(J,K)^T; the A operand T(I,K)[g][4kk+t] by the same intra-quad pattern
// Reports steps/s, the DMMA rate, and checks one filter against plain host loops.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
constexpr int NB = 5, N = 8 * NB, UPD = 2;
#ifndef WPC
#define WPC 4
#endif
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// element (g, 4kk + t) of an accumulator block held as (c0, c1): from lane 4g + 2kk + t/2, register t % 2
__device__ __forceinline__ double frag_elem(double c0, double c1, int src, bool odd) {
  const double x0 = __shfl_sync(0xffffffffu, c0, src), x1 = __shfl_sync(0xffffffffu, c1, src);
  return odd ? x1 : x0;
}
__global__ void __launch_bounds__(32 * WPC) frag_kernel(double* P, const double* Ag, const double* GK, int filters, int steps) {
  extern __shared__ double sm[];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  double* sA = sm + (size_t)w * (N * N + N * 4);  // A row-major [40][40]
  double* sGK = sA + N * N;                       // [40][4] = (g0, g1, k0, k1)
  const int src0 = (lane & ~3) + (t >> 1), src1 = src0 + 2;
  const bool odd = t & 1;
  for (int f = blockIdx.x * WPC + w; f < filters; f += gridDim.x * WPC) {
    double c0[NB][NB], c1[NB][NB];
    double* Pf = P + (size_t)f * NB * NB * 64;
#pragma unroll
    for (int I = 0; I < NB; I++)
#pragma unroll
      for (int J = 0; J < NB; J++) {
        const double2 v = *(const double2*)(Pf + ((I * NB + J) * 32 + lane) * 2);
        c0[I][J] = v.x;
        c1[I][J] = v.y;
      }
    for (int i = lane; i < N * N; i += 32) sA[i] = Ag[i];
    for (int i = lane; i < N * 4; i += 32) sGK[i] = GK[i];
    __syncwarp();
    for (int s = 0; s < steps; s++) {
      // ---- T = A P ----
      double t0[NB][NB], t1[NB][NB];
#pragma unroll
      for (int I = 0; I < NB; I++)
#pragma unroll
        for (int J = 0; J < NB; J++) t0[I][J] = t1[I][J] = 0.0;
#pragma unroll
      for (int K = 0; K < NB; K++)
#pragma unroll
        for (int kk = 0; kk < 2; kk++) {
          double a[NB];
#pragma unroll
          for (int I = 0; I < NB; I++) a[I] = sA[(8 * I + g) * N + 8 * K + 4 * kk + t];
#pragma unroll
          for (int J = 0; J < NB; J++) {
            const double b = frag_elem(c0[J][K], c1[J][K], kk ? src1 : src0, odd);  // P(K,J)[4kk+t][g] = P(J,K)[g][4kk+t]
#pragma unroll
            for (int I = 0; I < NB; I++) dmma(t0[I][J], t1[I][J], a[I], b);
          }
        }
      // ---- P' = T A^T (overwrites P) ----
#pragma unroll
      for (int I = 0; I < NB; I++)
#pragma unroll
        for (int J = 0; J < NB; J++) c0[I][J] = c1[I][J] = 0.0;
#pragma unroll
      for (int K = 0; K < NB; K++)
#pragma unroll
        for (int kk = 0; kk < 2; kk++) {
          double b[NB];
#pragma unroll
          for (int J = 0; J < NB; J++) b[J] = sA[(8 * J + g) * N + 8 * K + 4 * kk + t];  // A(J,K)^T[4kk+t][g]
#pragma unroll
          for (int I = 0; I < NB; I++) {
            const double a = frag_elem(t0[I][K], t1[I][K], kk ? src1 : src0, odd);
#pragma unroll
            for (int J = 0; J < NB; J++) dmma(c0[I][J], c1[I][J], a, b[J]);
          }
        }
      // ---- UPD x (P += G K^T + K G^T): A operand row (g0 g1 k0 k1), B operand rows (k0 k1 g0 g1) ----
      for (int u = 0; u < UPD; u++) {
        double a[NB], b[NB];
#pragma unroll
        for (int I = 0; I < NB; I++) {
          a[I] = sGK[(8 * I + g) * 4 + t];
          b[I] = sGK[(8 * I + g) * 4 + (t ^ 2)];
        }
#pragma unroll
        for (int I = 0; I < NB; I++)
#pragma unroll
          for (int J = 0; J < NB; J++) dmma(c0[I][J], c1[I][J], a[I], b[J]);
      }
    }
#pragma unroll
    for (int I = 0; I < NB; I++)
#pragma unroll
      for (int J = 0; J < NB; J++) *(double2*)(Pf + ((I * NB + J) * 32 + lane) * 2) = make_double2(c0[I][J], c1[I][J]);
    __syncwarp();
  }
}
int main() {
  cudaDeviceProp pr;
  cudaGetDeviceProperties(&pr, 0);
  const int steps = 20;
  const size_t smem = sizeof(double) * (N * N + N * 4) * WPC;
  cudaFuncSetAttribute(frag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, frag_kernel, 32 * WPC, smem);
  const int filters = pr.multiProcessorCount * occ * WPC * 8;
  std::vector<double> hA(N * N), hGK(N * 4), hP((size_t)N * N);
  srand(7);
  for (int i = 0; i < N; i++)
    for (int j = 0; j < N; j++) hA[i * N + j] = (i == j ? 1.0 : 0.0) + (rand() / (double)RAND_MAX - 0.5) * 0.01;  // I + small M
  for (auto& v : hGK) v = (rand() / (double)RAND_MAX - 0.5) * 0.05;
  for (int i = 0; i < N; i++)
    for (int j = 0; j <= i; j++) hP[i * N + j] = hP[j * N + i] = (i == j ? 1.0 : 0.0) + (rand() / (double)RAND_MAX - 0.5) * 0.02;
  // fragment-major copy of P for every filter
  std::vector<double> hF((size_t)NB * NB * 64);
  for (int I = 0; I < NB; I++)
    for (int J = 0; J < NB; J++)
      for (int l = 0; l < 32; l++)
        for (int e = 0; e < 2; e++) hF[((I * NB + J) * 32 + l) * 2 + e] = hP[(8 * I + (l >> 2)) * N + 8 * J + 2 * (l & 3) + e];
  double *dP, *dA, *dGK;
  cudaMalloc(&dP, sizeof(double) * hF.size() * (size_t)filters);
  cudaMalloc(&dA, sizeof(double) * hA.size());
  cudaMalloc(&dGK, sizeof(double) * hGK.size());
  cudaMemcpy(dA, hA.data(), sizeof(double) * hA.size(), cudaMemcpyHostToDevice);
  cudaMemcpy(dGK, hGK.data(), sizeof(double) * hGK.size(), cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float ms = 0;
  for (int rep = 0; rep < 3; rep++) {
    for (int f = 0; f < filters; f++) cudaMemcpyAsync(dP + (size_t)f * hF.size(), hF.data(), sizeof(double) * hF.size(), cudaMemcpyHostToDevice);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    frag_kernel<<<pr.multiProcessorCount * occ, 32 * WPC, smem>>>(dP, dA, dGK, filters, steps);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
  }
  printf("launch: %s, %d CTAs x %d warps per SM, %zu B shared memory per CTA\n", cudaGetErrorString(cudaGetLastError()), occ, WPC, smem);
  std::vector<double> got(hF.size());
  cudaMemcpy(got.data(), dP + (size_t)(filters - 1) * hF.size(), sizeof(double) * hF.size(), cudaMemcpyDeviceToHost);
  // host reference: the same sequence with plain loops
  std::vector<double> Pr = hP, T(N * N), Pn(N * N);
  for (int s = 0; s < steps; s++) {
    for (int i = 0; i < N; i++)
      for (int j = 0; j < N; j++) {
        double acc = 0;
        for (int k = 0; k < N; k++) acc = fma(hA[i * N + k], Pr[k * N + j], acc);
        T[i * N + j] = acc;
      }
    for (int i = 0; i < N; i++)
      for (int j = 0; j < N; j++) {
        double acc = 0;
        for (int k = 0; k < N; k++) acc = fma(T[i * N + k], hA[j * N + k], acc);
        Pn[i * N + j] = acc;
      }
    Pr = Pn;
    for (int u = 0; u < UPD; u++)
      for (int i = 0; i < N; i++)
        for (int j = 0; j < N; j++)
          Pr[i * N + j] += hGK[i * 4 + 0] * hGK[j * 4 + 2] + hGK[i * 4 + 1] * hGK[j * 4 + 3] + hGK[i * 4 + 2] * hGK[j * 4 + 0] + hGK[i * 4 + 3] * hGK[j * 4 + 1];
  }
  double worst = 0, asym = 0;
  for (int I = 0; I < NB; I++)
    for (int J = 0; J < NB; J++)
      for (int l = 0; l < 32; l++)
        for (int e = 0; e < 2; e++) {
          const int r = 8 * I + (l >> 2), c = 8 * J + 2 * (l & 3) + e;
          const double v = got[((I * NB + J) * 32 + l) * 2 + e];
          worst = fmax(worst, fabs(v - Pr[r * N + c]) / fmax(fabs(Pr[r * N + c]), 1e-3));
          const int l2 = (c % 8) * 4 + (r % 8) / 2, e2 = (r % 8) % 2;
          asym = fmax(asym, fabs(v - got[((J * NB + I) * 32 + l2) * 2 + e2]));
        }
  const double sps = (double)filters * steps / (ms * 1e-3);
  const double dmma_per_step = 2.0 * NB * NB * NB * 2 + UPD * NB * NB;
  printf("n = 40 covariance on DMMA fragments, one warp per filter: %.1f M (propagate + %d rank-4 updates)/s on %d SMs (%.2f ms, %d filters x %d steps)\n"
         "  %.0f DMMA per step -> %.2f TFLOP/s of DMMA work; worst rel err vs plain loops %.2e, worst |P - P^T| %.2e\n",
         sps * 1e-6, UPD, pr.multiProcessorCount, ms, filters, steps, dmma_per_step, sps * dmma_per_step * 512.0 * 1e-12, worst, asym);
  printf("for scale: the round-2 engine runs 19.9 M filter-steps/s at 65,536 x 8 (one structured propagate + ~2.3 updates per step)\n");
  return 0;
}
