/*
 * viekf_noise.h -- counter-based per-instance noise shared by the CUDA engine and the CPU oracle.
 *
 * The reference draws sensor noise from std::normal_distribution inside lib/multirotor_sim
 * (src/simulator.cpp:384-403 IMU, :406-488 camera); a batch of millions of Monte-Carlo instances cannot ship
 * one stream per instance through the ABI, so every instance regenerates its own stream from
 * Philox-4x32-10 keyed by (seed) with counter (instance, event, component pair) -- SURVEY.md 8(d).
 * The same header compiles for host (oracle, tests) and device (csrc/), so both sides draw bit-identical
 * 32-bit words; the double-precision Box-Muller on top differs only by libm-vs-libdevice rounding and by the device's exact
 * argument reduction (sincospi(2 u2) against cos / sin of the rounded product 2 pi u2): <= 2e-15 absolute on an N(0,1) draw.
 */
#ifndef VIEKF_NOISE_H
#define VIEKF_NOISE_H

#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define VIEKF_HD __host__ __device__ __forceinline__
#else
#define VIEKF_HD static inline
#endif

/* Philox-4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11). */
VIEKF_HD void viekf_philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                               uint32_t out[4]) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  for (int r = 0; r < 10; r++) {
    const uint64_t p0 = (uint64_t)M0 * c0, p1 = (uint64_t)M1 * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    const uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* Two independent N(0,1) draws for (seed, instance, event, pair). */
VIEKF_HD void viekf_normal_pair(uint64_t seed, uint64_t instance, uint32_t event, uint32_t pair, double n[2]) {
  uint32_t w[4];
  viekf_philox4x32((uint32_t)instance, (uint32_t)(instance >> 32), event, pair, (uint32_t)seed, (uint32_t)(seed >> 32), w);
  const double u1 = (double)((((uint64_t)w[0] << 21) ^ ((uint64_t)w[1] >> 11)) + 1ull) * (1.0 / 9007199254740992.0); /* (0,1] */
  const double u2 = (double)(((uint64_t)w[2] << 21) ^ ((uint64_t)w[3] >> 11)) * (1.0 / 9007199254740992.0);         /* [0,1) */
  const double r = sqrt(-2.0 * log(u1));
#if defined(__CUDA_ARCH__)
  double sn, cn; /* one exact argument reduction for both (sincospi); see the header comment for the distance to the host's values */
  sincospi(2.0 * u2, &sn, &cn);
  n[0] = r * cn;
  n[1] = r * sn;
#else
  const double a = 6.283185307179586476925286766559 * u2;
  n[0] = r * cos(a);
  n[1] = r * sin(a);
#endif
}

/* U(-1,1) per (seed, instance, k): constant per-instance quantities (IMU bias, initial-state perturbation use
 * event = 0xFFFFFFFF / 0xFFFFFFFE so they never collide with an event's stream). */
VIEKF_HD double viekf_uniform_pm1(uint64_t seed, uint64_t instance, uint32_t event, uint32_t k) {
  uint32_t w[4];
  viekf_philox4x32((uint32_t)instance, (uint32_t)(instance >> 32), event, k, (uint32_t)seed, (uint32_t)(seed >> 32), w);
  const double u = (double)(((uint64_t)w[0] << 21) ^ ((uint64_t)w[1] >> 11)) * (1.0 / 9007199254740992.0);
  return 2.0 * u - 1.0;
}

#define VIEKF_NOISE_EVENT_BIAS 0xFFFFFFFFu
#define VIEKF_NOISE_EVENT_X0 0xFFFFFFFEu

#endif
