// viekf_dev.cuh -- device-side building blocks shared by the tape executors (viekf_kernel.cu: covariance in shared
// memory; viekf_tile.cu: feature covariance in register tiles): shared-memory layout, per-filter context, dynamics and
// Jacobian blocks, manifold helpers, measurement models, S^-1 / gate, feature bookkeeping, statistics, value resolution.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "viekf_engine.h"
#include "viekf_math.cuh"
#include "../../include/viekf_noise.h"

namespace vk {

// ------------------------------------------------------------------------------------------------------------------
// shared-memory layout of one filter (offsets in doubles), a function of N only
// ------------------------------------------------------------------------------------------------------------------
struct Layout {
  int x, P, lam, dx, Abd, AbT, T, Ac, Gb, Afd, Gd, Y, upd, small, ids, total;
};
__host__ __device__ constexpr int ev2(int v) { return (v + 1) & ~1; }
__host__ __device__ inline Layout make_layout(int N) {
  Layout L;
  const int nmax = 16 + 3 * N;
  int o = 0;
  L.x = o;     o += ev2(17 + 5 * N);
  L.P = o;     o += ev2(nmax * (nmax + 1) / 2);
  L.lam = o;   o += ev2(nmax);
  L.dx = o;    o += ev2(nmax);
  L.Abd = o;   o += 256;
  L.AbT = o;   o += 256;  // A_bb transposed (bank-conflict-free column access in P'_fb = Y A_bb^T)
  L.T = o;     o += 256;
  L.Ac = o;    o += 144;
  L.Gb = o;    o += 54 + 2;
  L.Afd = o;   o += ev2(3 * N * 19);
  L.Gd = o;    o += ev2(nmax * 6);
  L.Y = o;     o += ev2(3 * N * 16);  // also the continuous feature blocks (27 per feature) before discretisation
  L.upd = L.Afd;                      // update workspace (P H^T, K, K S, 1-lambda: 10 n) overlays the propagate workspace
  const int upd_end = L.upd + ev2(10 * nmax);
  if (upd_end > o) o = upd_end;
  L.small = o; o += 80;
  L.ids = o;   o += ev2(N + 12);  // ids[N] + rm[N] + scalars[SC_COUNT + 9], as int32
  L.total = ev2(o);
  return L;
}
// scalars kept as int32 behind the ids array
enum { SC_LEN = 0, SC_NEXT = 1, SC_STARTED = 2, SC_FLAGS = 3, SC_NGATED = 4, SC_NFIX = 5, SC_NNAN = 6, SC_NKF = 7, SC_TMP0 = 8, SC_TMP1 = 9,
       SC_COUNT = 10 };

template <int G>
struct Ctx {
  double *x, *P, *lam, *dx, *Abd, *AbT, *T, *Ac, *Gb, *Afd, *Gd, *Y, *upd, *sm;
  int *ids, *rm, *sc;
  double* tcur;  // filter time lives in sm[SM_T]
  const double* zv;  // this event's resolved per-instance values (one of the two SM_ZV buffers)
  int tid, gi;
  int64_t b;
  __device__ __forceinline__ void sync() const {
    if (G == 32)
      __syncwarp();
    else
      asm volatile("bar.sync %0, %1;" ::"r"(gi + 1), "r"(G) : "memory");
  }
};

// small scratch slots inside sm[80]
enum { SM_ZV = 0 /*2 x 8, double buffered*/, SM_HC = 16 /*18*/, SM_RES = 34 /*4*/, SM_ER = 38 /*16*/, SM_UB = 54 /*6*/, SM_T = 60 };

// ------------------------------------------------------------------------------------------------------------------
// PROP
// ------------------------------------------------------------------------------------------------------------------
// Called by the whole last warp of the group: the zero fill of the continuous blocks is spread over its lanes, the
// (serial) dynamics and Jacobian entries are written by its last lane.
template <int G>
__device__ __forceinline__ void base_dynamics(const Ctx<G>& c, const KParams& p, const double* ub) {
  {
    const int lane = c.tid & 31;
    for (int i = lane; i < 144; i += 32) c.Ac[i] = 0.0;
    for (int i = lane; i < 54; i += 32) c.Gb[i] = 0.0;
    __syncwarp();
    if (lane != 31) return;
  }
  const double* x = c.x;
  V3 vel = ld3(x + 3);
  Q4 q = ldq(x + 6);
  V3 acc = mk3(ub[0] - x[10], ub[1] - x[11], ub[2] - x[12]);
  V3 om = mk3(ub[3] - x[13], ub[4] - x[14], ub[5] - x[15]);
  const double mu = x[16];
  double R[9];
  Rmat(q, R);
  V3 gB = rotp(q, mk3(0, 0, 9.80665));
  V3 pd = rota(q, vel);
  V3 oxv = cross(om, vel);
  V3 vd;
  if (p.use_drag)
    vd = mk3(gB.x - oxv.x - mu * vel.x, gB.y - oxv.y - mu * vel.y, acc.z + gB.z - oxv.z);
  else
    vd = acc + gB - oxv;
  double* dx = c.dx;
  dx[0] = pd.x; dx[1] = pd.y; dx[2] = pd.z;
  dx[3] = vd.x; dx[4] = vd.y; dx[5] = vd.z;
  dx[6] = om.x; dx[7] = om.y; dx[8] = om.z;
  for (int i = 9; i < 16; i++) dx[i] = 0.0;
  double* A = c.Ac;  // 9 x 16 row-major (rows POS, VEL, ATT of the continuous A), zero-filled above
  double* Gb = c.Gb;  // 9 x 6
  const double sv[9] = {0, -vel.z, vel.y, vel.z, 0, -vel.x, -vel.y, vel.x, 0};
  const double so[9] = {0, -om.z, om.y, om.z, 0, -om.x, -om.y, om.x, 0};
  const double sg[9] = {0, -gB.z, gB.y, gB.z, 0, -gB.x, -gB.y, gB.x, 0};
  for (int r = 0; r < 3; r++)
    for (int cc = 0; cc < 3; cc++) {
      A[r * 16 + 3 + cc] = R[cc * 3 + r];  // (POS,VEL) = R^T
      double s = 0;
      for (int k = 0; k < 3; k++) s += (-R[k * 3 + r]) * sv[k * 3 + cc];
      A[r * 16 + 6 + cc] = s;  // (POS,ATT) = -R^T [v]x
      A[(3 + r) * 16 + 3 + cc] = -so[r * 3 + cc];
      A[(3 + r) * 16 + 6 + cc] = sg[r * 3 + cc];
      A[(3 + r) * 16 + 12 + cc] = -sv[r * 3 + cc];
      A[(6 + r) * 16 + 6 + cc] = -so[r * 3 + cc];
      Gb[(3 + r) * 6 + 3 + cc] = -sv[r * 3 + cc];
    }
  for (int r = 0; r < 3; r++) {
    A[(6 + r) * 16 + 12 + r] = -1.0;
    Gb[(6 + r) * 6 + 3 + r] = -1.0;
  }
  if (p.use_drag) {
    A[3 * 16 + 3] += -mu;
    A[4 * 16 + 4] += -mu;
    A[5 * 16 + 11] = -1.0;
    A[3 * 16 + 15] = -vel.x;
    A[4 * 16 + 15] = -vel.y;
    Gb[5 * 6 + 2] = -1.0;
  } else {
    for (int r = 0; r < 3; r++) {
      A[(3 + r) * 16 + 9 + r] = -1.0;
      Gb[(3 + r) * 6 + r] = -1.0;
    }
  }
}

// continuous feature blocks of feature i: out[27] = A(f,VEL) 3x3 | A(f,B_G) 3x3 | A(f,f) 3x3, plus xdot (vi_ekf_dyn.cpp:96-134)
template <int G>
__device__ __forceinline__ void feature_dynamics(const Ctx<G>& c, const KParams& p, const double* ub, int i) {
  const double* x = c.x;
  V3 vel = ld3(x + 3);
  V3 om = mk3(ub[3] - x[13], ub[4] - x[14], ub[5] - x[15]);
  Q4 qbc = ldq(p.qbc);
  V3 pbc = ld3(p.pbc);
  V3 vc = rotp(qbc, vel + cross(om, pbc));
  V3 oc = rotp(qbc, om);
  Q4 qz = ldq(x + 17 + 5 * i);
  const double rho = x[17 + 5 * i + 4];
  const double rho2 = rho * rho;
  V3 ze, T0, T1;
  zeta_T(qz, ze, T0, T1);
  V3 zxv = cross(ze, vc);
  V3 w = oc + rho * zxv;
  double* dx = c.dx + 16 + 3 * i;
  dx[0] = -dot(T0, w);
  dx[1] = -dot(T1, w);
  dx[2] = rho2 * dot(ze, vc);
  double* o = c.Y + 27 * i;
  const double sz[9] = {0, -ze.z, ze.y, ze.z, 0, -ze.x, -ze.y, ze.x, 0};
  const double svc[9] = {0, -vc.z, vc.y, vc.z, 0, -vc.x, -vc.y, vc.x, 0};
  const double sw[9] = {0, -w.z, w.y, w.z, 0, -w.x, -w.y, w.x, 0};
  double M1[9];  // [zeta]x R_bc
  for (int r = 0; r < 3; r++)
    for (int cc = 0; cc < 3; cc++) {
      double s = 0;
      for (int k = 0; k < 3; k++) s += sz[r * 3 + k] * p.Rbc[k * 3 + cc];
      M1[r * 3 + cc] = s;
    }
  const double Tm[6] = {T0.x, T0.y, T0.z, T1.x, T1.y, T1.z};
  const double zv[3] = {ze.x, ze.y, ze.z};
  // M2 = rho [zeta]x R_bc [p_bc]x - R_bc, through the host-precomputed R_bc [p_bc]x
  double M2[9];
  for (int r = 0; r < 3; r++)
    for (int cc = 0; cc < 3; cc++) {
      double s = 0;
      for (int k = 0; k < 3; k++) s += sz[r * 3 + k] * p.RbcSp[k * 3 + cc];
      M2[r * 3 + cc] = rho * s - p.Rbc[r * 3 + cc];
    }
  double Mzz[9];
  for (int r = 0; r < 3; r++)
    for (int cc = 0; cc < 3; cc++) {
      double s = 0;
      for (int k = 0; k < 3; k++) s += svc[r * 3 + k] * sz[k * 3 + cc];
      Mzz[r * 3 + cc] = sw[r * 3 + cc] + rho * s;
    }
  for (int a = 0; a < 2; a++) {
    const double* Ta = Tm + 3 * a;
    for (int cc = 0; cc < 3; cc++) {
      o[a * 3 + cc] = -rho * (Ta[0] * M1[cc] + Ta[1] * M1[3 + cc] + Ta[2] * M1[6 + cc]);
      o[9 + a * 3 + cc] = -(Ta[0] * M2[cc] + Ta[1] * M2[3 + cc] + Ta[2] * M2[6 + cc]);
    }
    double tm[3];
    for (int cc = 0; cc < 3; cc++) tm[cc] = Ta[0] * Mzz[cc] + Ta[1] * Mzz[3 + cc] + Ta[2] * Mzz[6 + cc];
    for (int bb = 0; bb < 2; bb++) o[18 + a * 3 + bb] = -(tm[0] * Tm[3 * bb] + tm[1] * Tm[3 * bb + 1] + tm[2] * Tm[3 * bb + 2]);
    o[18 + a * 3 + 2] = -(Ta[0] * zxv.x + Ta[1] * zxv.y + Ta[2] * zxv.z);
  }
  for (int cc = 0; cc < 3; cc++) {
    o[6 + cc] = rho2 * (zv[0] * p.Rbc[cc] + zv[1] * p.Rbc[3 + cc] + zv[2] * p.Rbc[6 + cc]);
    o[15 + cc] = rho2 * (zv[0] * p.RbcSp[cc] + zv[1] * p.RbcSp[3 + cc] + zv[2] * p.RbcSp[6 + cc]);
  }
  double zs[3];
  for (int cc = 0; cc < 3; cc++) zs[cc] = zv[0] * svc[cc] + zv[1] * svc[3 + cc] + zv[2] * svc[6 + cc];
  for (int bb = 0; bb < 2; bb++) o[24 + bb] = rho2 * (zs[0] * Tm[3 * bb] + zs[1] * Tm[3 * bb + 1] + zs[2] * Tm[3 * bb + 2]);
  o[26] = 2 * rho * dot(ze, vc);
}

// fix_depth (vi_ekf_helper.cpp:128-156) for feature i
template <int G>
__device__ __forceinline__ void fix_depth_one(const Ctx<G>& c, const KParams& p, int i) {
  double* rho = c.x + 17 + 5 * i + 4;
  const int d = 16 + 3 * i + 2;
  const double reset = p.depth_reset;
  double r = *rho;
  uint32_t fl = 0;
  int nfix = 0;
  if (r != r) {
    r = reset;
    fl |= VIEKF_FLAG_DEPTH_NAN_FIXED;
    nfix++;
  }
  if (r < 0.0) {
    const double err = reset - r;
    c.P[plo(d, d)] += err * err;
    r = reset;
    fl |= VIEKF_FLAG_DEPTH_NEG_FIXED;
    nfix++;
  } else if (r > 1e2) {
    c.P[plo(d, d)] = p.P0f[2];
    r = reset;
    fl |= VIEKF_FLAG_DEPTH_BIG_FIXED;
    nfix++;
  }
  if (nfix) {
    *rho = r;
    atomicOr((unsigned*)&c.sc[SC_FLAGS], fl);
    atomicAdd(&c.sc[SC_NFIX], nfix);
  }
}

// x <- x [+] d  (vi_ekf_helper.cpp:88-98): base part by thread `base_tid` of the group, feature i by thread i.
// One out-of-line copy serves propagate and update (instruction-cache footprint).
static __device__ __noinline__ void boxplus_impl(double* x, const double* d, double s, int len, int tid, int G, int base_tid) {
  if (tid == base_tid) {
    for (int k = 0; k < 6; k++) x[k] += d[k] * s;
    Q4 q = otimes(ldq(x + 6), qexp(mk3(d[6] * s, d[7] * s, d[8] * s)));
    stq(x + 6, q);
    for (int k = 0; k < 7; k++) x[10 + k] += d[9 + k] * s;
  }
  for (int i = tid; i < len; i += G) {
    double* xf = x + 17 + 5 * i;
    const double* df = d + 16 + 3 * i;
    Q4 q = qfeat_boxplus(ldq(xf), df[0] * s, df[1] * s);
    stq(xf, q);
    xf[4] += df[2] * s;
  }
}
template <int G>
__device__ __forceinline__ void boxplus_inplace(const Ctx<G>& c, const double* d, double s, int len, int base_tid = G - 1) {
  boxplus_impl(c.x, d, s, len, c.tid, G, base_tid);
}

// Phase B of propagate_state: the discretised rows [A_d | G_d] = [I + A dt + A^2 dt^2/2 | (I + A dt/2 + A^2 dt^2/6) G dt]
// (vi_ekf.cpp:302-303) from the continuous blocks written by base_dynamics / feature_dynamics; one row per task, tasks
// first, first + stride, ... < npad; rows n..npad-1 (padding of the register-tile executor) are exact zeros.
// Every row is evaluated as "load a chunk of inputs into registers, compute, store": inputs and outputs live in the same
// shared-memory window, so with loads and stores interleaved the compiler has to assume aliasing and serialises one
// shared-memory round trip per output (this phase cost 8 k cycles per propagate that way).
template <int G>
__device__ __forceinline__ void discretise_rows(const Ctx<G>& c, double dt, int n, int npad, int first, int stride) {
  const double h2 = dt * dt / 2.0, h6 = dt * dt / 6.0, hd2 = dt / 2.0;
  for (int task = first; task < npad; task += stride) {
    if (task >= n) {  // padding rows of the register-tile executor: exact zeros
      double* Ad = c.Afd + (task - 16) * 19;
#pragma unroll
      for (int k = 0; k < 19; k++) Ad[k] = 0.0;
#pragma unroll
      for (int q = 0; q < 6; q++) c.Gd[task * 6 + q] = 0.0;
    } else if (task < 16) {
      const int r = task;
      double* Ad = c.Abd + r * 16;
      double* Gd = c.Gd + r * 6;
      if (r >= 9) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
          const double v = (k == r) ? 1.0 : 0.0;
          Ad[k] = v;
          c.AbT[k * 16 + r] = v;
        }
#pragma unroll
        for (int q = 0; q < 6; q++) Gd[q] = 0.0;
      } else {
        const double* Ar = c.Ac + r * 16;
        double ar[16];
#pragma unroll
        for (int k = 0; k < 16; k++) ar[k] = Ar[k];
        double m[9];
#pragma unroll
        for (int c0 = 0; c0 < 16; c0 += 4) {  // four columns at a time: 24 operands in registers, four independent chains
          double av[6][4];
#pragma unroll
          for (int k = 0; k < 6; k++)
#pragma unroll
            for (int j = 0; j < 4; j++) av[k][j] = c.Ac[(3 + k) * 16 + c0 + j];
          double o[4];
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const int cc = c0 + j;
            double aa = 0;
#pragma unroll
            for (int k = 3; k < 9; k++) aa += ar[k] * av[k - 3][j];
            const double id = (cc == r) ? 1.0 : 0.0;
            o[j] = (id + ar[cc] * dt) + aa * h2;
            if (cc < 9) m[cc] = (id + ar[cc] * hd2) + aa * h6;
          }
#pragma unroll
          for (int j = 0; j < 4; j++) {
            Ad[c0 + j] = o[j];
            c.AbT[(c0 + j) * 16 + r] = o[j];
          }
        }
        double gd[6];
#pragma unroll
        for (int q = 0; q < 6; q++) {
          double s = 0;
#pragma unroll
          for (int k = 0; k < 9; k++) s += m[k] * c.Gb[k * 6 + q];
          gd[q] = s * dt;
        }
#pragma unroll
        for (int q = 0; q < 6; q++) Gd[q] = gd[q];
      }
    } else {
      const int fr = task - 16, i = fr / 3, a = fr - 3 * i;
      const double* F = c.Y + 27 * i;  // Afv | Afg | Aff (continuous)
      double* Ad = c.Afd + fr * 19;
      double* Gd = c.Gd + task * 6;
      double fv[3], fg[3], ff[3];  // row a of the three continuous blocks
#pragma unroll
      for (int k = 0; k < 3; k++) {
        fv[k] = F[a * 3 + k];
        fg[k] = F[9 + a * 3 + k];
        ff[k] = F[18 + a * 3 + k];
      }
      double mfb[9];
#pragma unroll
      for (int cc = 0; cc < 3; cc++) mfb[cc] = 0.0;
      // columns VEL (3..5), ATT / B_A (6..11), B_G (12..14), MU (15), in three chunks of the VEL rows of the continuous A
#pragma unroll
      for (int c0 = 3; c0 < 16; c0 += 5) {
        double av[3][5], bk[3][3];
#pragma unroll
        for (int k = 0; k < 3; k++)
#pragma unroll
          for (int j = 0; j < 5; j++)
            if (c0 + j < 16) av[k][j] = c.Ac[(3 + k) * 16 + c0 + j];
        if (c0 == 3) {  // A(f,VEL) block: columns 3..5
#pragma unroll
          for (int k = 0; k < 9; k++) bk[k / 3][k % 3] = F[k];
        } else {        // A(f,B_G) block: columns 12..14
#pragma unroll
          for (int k = 0; k < 9; k++) bk[k / 3][k % 3] = F[9 + k];
        }
        double o[5];
#pragma unroll
        for (int j = 0; j < 5; j++) {
          const int cc = c0 + j;
          if (cc < 16) {
            double aa = 0;
#pragma unroll
            for (int k = 0; k < 3; k++) aa += fv[k] * av[k][j];
            double ac = 0.0;
            if (cc < 6) {
#pragma unroll
              for (int k = 0; k < 3; k++) aa += ff[k] * bk[k][cc - 3];
              ac = fv[cc - 3];
            } else if (cc >= 12 && cc < 15) {
#pragma unroll
              for (int k = 0; k < 3; k++) aa += ff[k] * bk[k][cc - 12];
              ac = fg[cc - 12];
            }
            o[j] = ac * dt + aa * h2;
            if (cc < 9) mfb[cc] = ac * hd2 + aa * h6;
          }
        }
        if (c0 == 3) {
#pragma unroll
          for (int cc = 0; cc < 3; cc++) Ad[cc] = 0.0;
        }
#pragma unroll
        for (int j = 0; j < 5; j++)
          if (c0 + j < 16) Ad[c0 + j] = o[j];
      }
      double mff[3];
      {
        double Bf[9];
#pragma unroll
        for (int k = 0; k < 9; k++) Bf[k] = F[18 + k];
        double o[3];
#pragma unroll
        for (int bb = 0; bb < 3; bb++) {
          double aa = 0;
#pragma unroll
          for (int k = 0; k < 3; k++) aa += ff[k] * Bf[k * 3 + bb];
          const double id = (a == bb) ? 1.0 : 0.0;
          o[bb] = (id + ff[bb] * dt) + aa * h2;
          mff[bb] = (id + ff[bb] * hd2) + aa * h6;
        }
#pragma unroll
        for (int bb = 0; bb < 3; bb++) Ad[16 + bb] = o[bb];
      }
      {
        double Bg[9], gd[6];
#pragma unroll
        for (int k = 0; k < 9; k++) Bg[k] = F[9 + k];
#pragma unroll
        for (int q = 0; q < 6; q++) {
          double s = 0;
#pragma unroll
          for (int k = 3; k < 9; k++) s += mfb[k] * c.Gb[k * 6 + q];
          if (q >= 3) {
#pragma unroll
            for (int k = 0; k < 3; k++) s += mff[k] * Bg[k * 3 + (q - 3)];
          }
          gd[q] = s * dt;
        }
#pragma unroll
        for (int q = 0; q < 6; q++) Gd[q] = gd[q];
      }
    }
  }
}

// Guards of propagate_state, continuous dynamics and the discretised rows [A_d | G_d] (phases A, B).  Returns false
// (uniformly) when the call does not advance the filter; `npad` >= n rows are written, rows n..npad-1 as zeros.
template <int G>
__device__ __forceinline__ bool prop_discretise(Ctx<G>& c, const KParams& p, const viekf_event& E, double& dt_out, int& len_out, int pad_to_even) {
  // the body-frame input of the latest propagate_state call is kept for UPD events with aux == 1 (vi_ekf.cpp:265-267)
  if (c.tid == 0) {
    const double* zz = c.zv;
    Q4 qbu = ldq(p.qbu);
    V3 a = rota(qbu, mk3(zz[0], zz[1], zz[2])), g = rota(qbu, mk3(zz[3], zz[4], zz[5]));
    double* ul = c.sm + SM_UB;
    ul[0] = a.x; ul[1] = a.y; ul[2] = a.z; ul[3] = g.x; ul[4] = g.y; ul[5] = g.z;
  }
  // uniform control flow: every thread evaluates the same guards from shared state (vi_ekf.cpp:274-289)
  const int started = c.sc[SC_STARTED];
  const double tcur = *c.tcur;
  const double dt = E.t - tcur;
  if (!started) {
    c.sync();
    if (c.tid == 0) {
      c.sc[SC_STARTED] = 1;
      *c.tcur = E.t;
    }
    c.sync();
    return false;
  }
  if (fabs(dt) < 1e-6) {
    c.sync();
    return false;
  }
  if (dt < 0) {
    if (c.tid == 0) c.sc[SC_FLAGS] |= VIEKF_FLAG_BACKWARD_DT;
    c.sync();
    return false;
  }
  const int len = c.sc[SC_LEN];
  const int n = 16 + 3 * len;
  const double* zv = c.zv;
  // IMU sample into the body frame (vi_ekf.cpp:265-267)
  double ub[6];
  {
    Q4 qbu = ldq(p.qbu);
    V3 a = rota(qbu, mk3(zv[0], zv[1], zv[2])), g = rota(qbu, mk3(zv[3], zv[4], zv[5]));
    ub[0] = a.x; ub[1] = a.y; ub[2] = a.z; ub[3] = g.x; ub[4] = g.y; ub[5] = g.z;
  }
  // ---- phase A: continuous dynamics and Jacobian blocks ----
  if (c.tid >= G - 32) base_dynamics(c, p, ub);
  for (int i = c.tid; i < len; i += G) feature_dynamics(c, p, ub, i);
  c.sync();
  // ---- phase B: discretisation, one row of [A_d | G_d] per task (vi_ekf.cpp:302-303) ----
  const int npad = pad_to_even ? 16 + 3 * ((len + 1) & ~1) : n;
  discretise_rows(c, dt, n, npad, c.tid, G);
  c.sync();
  dt_out = dt;
  len_out = len;
  return true;
}

template <int G>
__device__ void ev_propagate(Ctx<G>& c, const KParams& p, const viekf_event& E) {
  double dt;
  int len;
  if (!prop_discretise(c, p, E, dt, len, 0)) return;
  const int n = 16 + 3 * len;
  // ---- phase C: Y = A_fb P_bb + A_ff P_fb (3len x 16) and T = A_bb P_bb (16 x 16); reads the old P only ----
  const double* P = c.P;
  for (int task = c.tid; task < 48 * len + 256; task += G) {
    if (task < 48 * len) {
      const int fr = task >> 4, cc = task & 15, i = fr / 3;
      const double* Ad = c.Afd + fr * 19;
      double s = 0;
#pragma unroll
      for (int k = 3; k < 16; k++) s += Ad[k] * P[psym(k, cc)];
      const int f0 = 16 + 3 * i;
#pragma unroll
      for (int k = 0; k < 3; k++) s += Ad[16 + k] * P[plo(f0 + k, cc)];
      c.Y[task] = s;
    } else {
      const int t2 = task - 48 * len, r = t2 >> 4, cc = t2 & 15;
      double s = 0;
#pragma unroll
      for (int k = 0; k < 16; k++) s += c.Abd[r * 16 + k] * P[psym(k, cc)];
      c.T[t2] = s;
    }
  }
  c.sync();
  // ---- phase D: P'_ff blocks in place, and P'_bb from T ----
  const int nb = (len * (len + 1)) >> 1;
  for (int task = c.tid; task < nb + 136; task += G) {
    if (task < nb) {
      int i = (int)((sqrtf(8.0f * (float)task + 1.0f) - 1.0f) * 0.5f);
      while (((i + 1) * (i + 2) >> 1) <= task) i++;
      while (((i * (i + 1)) >> 1) > task) i--;
      const int j = task - ((i * (i + 1)) >> 1);
      const int fi = 16 + 3 * i, fj = 16 + 3 * j;
      const double* Ai = c.Afd + 3 * i * 19;
      const double* Aj = c.Afd + 3 * j * 19;
      const double* Yi = c.Y + 3 * i * 16;
      double Z[9];
#pragma unroll
      for (int a = 0; a < 3; a++)
#pragma unroll
        for (int bb = 0; bb < 3; bb++) {
          double s = 0;
          const double* Pr = c.P + plo(fj + bb, 0);
#pragma unroll
          for (int k = 3; k < 16; k++) s += Ai[a * 19 + k] * Pr[k];
#pragma unroll
          for (int k = 0; k < 3; k++) s += Ai[a * 19 + 16 + k] * c.P[psym(fi + k, fj + bb)];
          Z[a * 3 + bb] = s;
        }
      double o[9];
#pragma unroll
      for (int a = 0; a < 3; a++)
#pragma unroll
        for (int bb = 0; bb < 3; bb++) {
          double s = 0;
#pragma unroll
          for (int k = 3; k < 16; k++) s += Yi[a * 16 + k] * Aj[bb * 19 + k];
#pragma unroll
          for (int k = 0; k < 3; k++) s += Z[a * 3 + k] * Aj[bb * 19 + 16 + k];
#pragma unroll
          for (int q = 0; q < 6; q++) s += (c.Gd[(fi + a) * 6 + q] * p.Qu[q]) * c.Gd[(fj + bb) * 6 + q];
          o[a * 3 + bb] = s;
        }
      if (i == j) {
        for (int a = 0; a < 3; a++)
          for (int bb = 0; bb <= a; bb++) c.P[plo(fi + a, fj + bb)] = o[a * 3 + bb] + (a == bb ? p.Qxf[a] : 0.0);
      } else {
        for (int a = 0; a < 3; a++)
          for (int bb = 0; bb < 3; bb++) c.P[plo(fi + a, fj + bb)] = o[a * 3 + bb];
      }
    } else {
      const int e = task - nb;
      int r = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
      while (((r + 1) * (r + 2) >> 1) <= e) r++;
      while (((r * (r + 1)) >> 1) > e) r--;
      const int cc = e - ((r * (r + 1)) >> 1);
      double s = 0;
#pragma unroll
      for (int k = 0; k < 16; k++) s += c.T[r * 16 + k] * c.Abd[cc * 16 + k];
#pragma unroll
      for (int q = 0; q < 6; q++) s += (c.Gd[r * 6 + q] * p.Qu[q]) * c.Gd[cc * 6 + q];
      if (r == cc) s += p.Qx[r];
      c.P[e] = s;
    }
  }
  c.sync();
  // ---- phase E: P'_fb = Y A_bb^T (+ G_d Qu G_d^T), state boxplus ----
  for (int task = c.tid; task < 48 * len; task += G) {
    const int fr = task >> 4, cc = task & 15;
    double s = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) s += c.Y[fr * 16 + k] * c.AbT[k * 16 + cc];
#pragma unroll
    for (int q = 0; q < 6; q++) s += (c.Gd[(16 + fr) * 6 + q] * p.Qu[q]) * c.Gd[cc * 6 + q];
    c.P[plo(16 + fr, cc)] = s;
  }
  boxplus_inplace(c, c.dx, dt, len);
  if (c.tid == 0) *c.tcur = E.t;
  c.sync();
  // ---- phase F: fix_depth ----
  for (int i = c.tid; i < len; i += G) fix_depth_one(c, p, i);
  c.sync();
}

// ------------------------------------------------------------------------------------------------------------------
// feature management
// ------------------------------------------------------------------------------------------------------------------
template <int G>
__device__ __forceinline__ int find_feature(const Ctx<G>& c, int id) {
  // global_to_local_feature_id (vi_ekf_helper.cpp:114-125): every warp searches the id list with one ballot per 32 ids
  const int len = c.sc[SC_LEN];
  const int lane = c.tid & 31;
  for (int base = 0; base < len; base += 32) {
    const int i = base + lane;
    const unsigned hit = __ballot_sync(0xffffffffu, i < len && c.ids[i] == id);
    if (hit) return base + __ffs(hit) - 1;
  }
  return -1;
}

// init_feature (vi_ekf_feat.cpp:6-47).  Returns (uniformly) whether a slot was available.
template <int G>
__device__ bool do_init_feature(Ctx<G>& c, const KParams& p, double l0, double l1, double depth) {
  const int len = c.sc[SC_LEN];
  if (len >= p.N) return false;
  const int r0 = 16 + 3 * len;
  c.sync();
  if (c.tid == 0) {
    const double lc0 = l0 - p.cc[0], lc1 = l1 - p.cc[1];
    V3 z = mk3(lc0, lc1 * (p.fl[1] / p.fl[0]), p.fl[0]);
    const double nz = norm(z);
    z = mk3(z.x / nz, z.y / nz, z.z / nz);
    Q4 q = from_two_unit_vectors(mk3(0, 0, 1), z);
    double d = depth;
    if (d != d) d = 2.0 * p.min_depth;
    double* xf = c.x + 17 + 5 * len;
    stq(xf, q);
    xf[4] = 1.0 / d;
    c.ids[len] = c.sc[SC_NEXT];
    c.sc[SC_NEXT] += 1;
    c.sc[SC_LEN] = len + 1;
  }
  double* Pn = c.P + plo(r0, 0);
  const int cnt = plo(r0 + 3, 0) - plo(r0, 0);
  for (int e = c.tid; e < cnt; e += G) Pn[e] = 0.0;
  c.sync();
  if (c.tid < 3) c.P[plo(r0 + c.tid, r0 + c.tid)] = p.P0f[c.tid];
  c.sync();
  return true;
}

// clear_feature (vi_ekf_feat.cpp:50-73): order-preserving compaction of x and of the packed P, in place.
template <int G>
__device__ void do_clear_feature(Ctx<G>& c, const KParams& p, int local) {
  const int len = c.sc[SC_LEN];
  if (local < 0 || local >= len) return;
  c.sync();
  const int nl = len - 1;
  // x tail
  {
    const int dst0 = 17 + 5 * local, cnt = 5 * (nl - local);
    for (int base = 0; base < cnt; base += G) {
      const int e = base + c.tid;
      double v = 0;
      if (e < cnt) v = c.x[dst0 + 5 + e];
      c.sync();
      if (e < cnt) c.x[dst0 + e] = v;
      c.sync();
    }
  }
  // P: new(i,j) = old(i + 3[i>=d0], j + 3[j>=d0]); destination index <= source index, so chunks in ascending order with
  // a read-all / barrier / write-all cycle never overwrite an element that is still to be read.
  const int d0 = 16 + 3 * local;
  const int nn = 16 + 3 * nl;
  const int first = plo(d0, 0), last = plo(nn, 0);
  constexpr int PER = 4;
  for (int base = first; base < last; base += G * PER) {
    double v[PER];
#pragma unroll
    for (int u = 0; u < PER; u++) {
      const int e = base + u * G + c.tid;
      v[u] = 0;
      if (e < last) {
        int i = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
        while (((i + 1) * (i + 2) >> 1) <= e) i++;
        while (((i * (i + 1)) >> 1) > e) i--;
        const int j = e - ((i * (i + 1)) >> 1);
        v[u] = c.P[plo(i + 3, j >= d0 ? j + 3 : j)];
      }
    }
    c.sync();
#pragma unroll
    for (int u = 0; u < PER; u++) {
      const int e = base + u * G + c.tid;
      if (e < last) c.P[e] = v[u];
    }
    c.sync();
  }
  if (c.tid == 0) {
    for (int i = local; i < nl; i++) c.ids[i] = c.ids[i + 1];
    c.sc[SC_LEN] = nl;
  }
  c.sync();
}

// Thread-0 part of keyframe_reset (vi_ekf_kfr.cpp:56-120): zero position, strip yaw, the 3x3 attitude block Nb of N, and the
// global pose / covariance accumulation.  Reads only the base block of c.P (packed lower triangle, indices < 136).
template <int G>
__device__ void kfreset_pose(Ctx<G>& c, const KParams& p, double* Nb) {
  {
    FilterMeta& M = p.meta[c.b];
    double* x = c.x;
    double ecov[36];
    for (int k = 0; k < 36; k++) ecov[k] = 0.0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) ecov[i + 6 * j] = c.P[psym(i, j)];
    ecov[5 + 6 * 5] = c.P[plo(8, 8)];
    V3 et = ld3(x);
    x[0] = x[1] = x[2] = 0.0;
    Q4 q = ldq(x + 6);
    const double yaw = qyaw(q), roll = qroll(q), pitch = qpitch(q);
    Q4 qe = from_euler(0, 0, yaw);
    stq(x + 6, from_euler(roll, pitch, 0.0));
    const double cp = cos(roll), sp = sin(roll), tt = tan(pitch);
    Nb[0] = 1; Nb[1] = sp * tt; Nb[2] = cp * tt;
    Nb[3] = 0; Nb[4] = cp * cp; Nb[5] = -cp * sp;
    Nb[6] = 0; Nb[7] = -cp * sp; Nb[8] = sp * sp;
    // global covariance: P_new = P_prev + Adj^T cov Adj, Adj = [[R,[t]x R],[0,R]] of the current node pose
    Q4 nq = ldq(M.node_q);
    V3 nt = ld3(M.node_t);
    double R[9];
    Rmat(nq, R);
    const double st[9] = {0, -nt.z, nt.y, nt.z, 0, -nt.x, -nt.y, nt.x, 0};
    double Adj[36];
    for (int k = 0; k < 36; k++) Adj[k] = 0.0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += st[i * 3 + k] * R[k * 3 + j];
        Adj[i + 6 * j] = R[i * 3 + j];
        Adj[i + 6 * (j + 3)] = s;
        Adj[(i + 3) + 6 * (j + 3)] = R[i * 3 + j];
      }
    double Tm[36];
    for (int i = 0; i < 6; i++)
      for (int j = 0; j < 6; j++) {
        double s = 0;
        for (int k = 0; k < 6; k++) s += Adj[k + 6 * i] * ecov[k + 6 * j];
        Tm[i + 6 * j] = s;
      }
    for (int i = 0; i < 6; i++)
      for (int j = 0; j < 6; j++) {
        double s = 0;
        for (int k = 0; k < 6; k++) s += Tm[i + 6 * k] * Adj[k + 6 * j];
        M.gcov[i + 6 * j] += s;
      }
    V3 nt2 = nt + rota(nq, et);
    M.node_t[0] = nt2.x; M.node_t[1] = nt2.y; M.node_t[2] = nt2.z;
    stq(M.node_q, otimes(nq, qe));
  }
}

// keyframe_reset (vi_ekf_kfr.cpp:56-157): zero position, strip yaw, P = N P N^T, accumulate the global pose / cov.
template <int G>
__device__ void do_keyframe_reset(Ctx<G>& c, const KParams& p) {
  const int len = c.sc[SC_LEN];
  const int n = 16 + 3 * len;
  c.sync();
  double* Nb = c.sm + SM_HC;  // 9 doubles, row-major
  if (c.tid == 0) kfreset_pose(c, p, Nb);
  c.sync();
  // P = N P N^T with N = I except N(POS,POS) = 0 and N(ATT,ATT) = Nb
  for (int i = c.tid; i < n; i += G) {
    if (i < 3) {
      for (int j = 0; j <= i; j++) c.P[plo(i, j)] = 0.0;
    } else if (i < 6) {
      for (int j = 0; j < 3; j++) c.P[plo(i, j)] = 0.0;
      // column part: rows 6..8 against this column i (i in 3..5)
      double v[3], o[3];
      for (int a = 0; a < 3; a++) v[a] = c.P[plo(6 + a, i)];
      for (int a = 0; a < 3; a++) o[a] = Nb[a * 3] * v[0] + Nb[a * 3 + 1] * v[1] + Nb[a * 3 + 2] * v[2];
      for (int a = 0; a < 3; a++) c.P[plo(6 + a, i)] = o[a];
    } else if (i == 6) {
      for (int r = 6; r < 9; r++)
        for (int j = 0; j < 3; j++) c.P[plo(r, j)] = 0.0;
      double A[9], B[9];
      for (int a = 0; a < 3; a++)
        for (int bb = 0; bb < 3; bb++) A[a * 3 + bb] = c.P[psym(6 + a, 6 + bb)];
      for (int a = 0; a < 3; a++)
        for (int bb = 0; bb < 3; bb++) B[a * 3 + bb] = Nb[a * 3] * A[bb] + Nb[a * 3 + 1] * A[3 + bb] + Nb[a * 3 + 2] * A[6 + bb];
      for (int a = 0; a < 3; a++)
        for (int bb = 0; bb <= a; bb++)
          c.P[plo(6 + a, 6 + bb)] = B[a * 3] * Nb[bb * 3] + B[a * 3 + 1] * Nb[bb * 3 + 1] + B[a * 3 + 2] * Nb[bb * 3 + 2];
    } else if (i >= 9) {
      for (int j = 0; j < 3; j++) c.P[plo(i, j)] = 0.0;
      double v[3];
      for (int a = 0; a < 3; a++) v[a] = c.P[plo(i, 6 + a)];
      for (int a = 0; a < 3; a++) c.P[plo(i, 6 + a)] = Nb[a * 3] * v[0] + Nb[a * 3 + 1] * v[1] + Nb[a * 3 + 2] * v[2];
    }
  }
  c.sync();
}

// keep_only_features (vi_ekf_feat.cpp:81-142)
template <int G>
__device__ void do_keep_only(Ctx<G>& c, const KParams& p, const int32_t* feats, int nf) {
  c.sync();
  // one thread per tracked feature decides keep / drop and keyframe overlap; thread 0 then lists the drops in order
  int* keepf = (int*)c.upd;  // [N] flags, update workspace is idle here
  if (c.tid == 0) c.sc[SC_TMP1] = 0;
  c.sync();
  {
    const int len = c.sc[SC_LEN];
    const int nkf = c.sc[SC_NKF];
    const int32_t* kf = p.kf_ids + c.b * KF_CAP;
    for (int l = c.tid; l < len; l += G) {
      const int id = c.ids[l];
      bool keep = false;
      for (int i = 0; i < nf; i++) keep |= (feats[i] == id);
      if (keep && p.use_kfreset) {
        bool ov = false;
        for (int j = 0; j < nkf; j++) ov |= (kf[j] == id);
        if (ov) atomicAdd(&c.sc[SC_TMP1], 1);
      }
      keepf[l] = keep ? 1 : 0;
    }
  }
  c.sync();
  if (c.tid == 0) {
    const int len = c.sc[SC_LEN];
    int nrm = 0;
    for (int l = 0; l < len; l++)
      if (!keepf[l]) c.rm[nrm++] = c.ids[l];
    c.sc[SC_TMP0] = nrm;
  }
  c.sync();
  const int nrm = c.sc[SC_TMP0];
  const int overlap = c.sc[SC_TMP1];
  const int nkf = c.sc[SC_NKF];
  for (int k = 0; k < nrm; k++) do_clear_feature(c, p, find_feature(c, c.rm[k]));
  const bool reset = p.use_kfreset && nkf > 0 && ((double)overlap / (double)nkf < p.kf_thresh);
  if (reset) do_keyframe_reset(c, p);
  c.sync();  // every thread has taken the (uniform) reset decision from the old list size before thread 0 replaces it
  if (p.use_kfreset && (reset || nkf == 0)) {
    int32_t* kf = p.kf_ids + c.b * KF_CAP;
    const int m = nf < KF_CAP ? nf : KF_CAP;
    for (int i = c.tid; i < m; i += G) kf[i] = feats[i];
    if (c.tid == 0) c.sc[SC_NKF] = m;
  }
  c.sync();
}

// ------------------------------------------------------------------------------------------------------------------
// UPD
// ------------------------------------------------------------------------------------------------------------------
// symmetric rank update of the packed P:
//   P_ij += (1 - a_i a_j) * ( KS_i.K_j - K_i.H_j - H_i.K_j ),   H = (HP)^T = P H^T,   a = 1 - lambda (partial update) or 0
// Work item = 4 rows x all 32-column chunks of that row block; the row vectors are loaded once per row block
// (broadcast), the column vectors once per chunk (16-byte vector loads, conflict free), P is read and written once.
// Vectors live in the update workspace: Kv/Hv/KSv hold components 0,1 as double2, K2/H2/KS2 component 2 (RD == 3).
template <int G>
struct UpdVec {
  double2 *Kv, *Hv, *KSv;
  double *K2, *H2, *KS2, *av;
  __device__ __forceinline__ UpdVec(const Ctx<G>& c, int nmax) {
    double* w = c.upd;
    Kv = (double2*)w;
    Hv = (double2*)(w + 2 * nmax);
    KSv = (double2*)(w + 4 * nmax);
    K2 = w + 6 * nmax;
    H2 = w + 7 * nmax;
    KS2 = w + 8 * nmax;
    av = w + 9 * nmax;
  }
};

template <int G, int RD>
__device__ __noinline__ void rank_update(const Ctx<G>& c, const UpdVec<G>& v, int n) {
  const int lane = c.tid & 31, warp = c.tid >> 5;
  constexpr int NW = G / 32;
  constexpr int RB = 4;
  const int nrb = (n + RB - 1) / RB;
  for (int rb = warp; rb < nrb; rb += NW) {
    const int i0 = rb * RB;
    double2 ki[RB], hi[RB], si[RB];
    double k2[RB], h2[RB], s2[RB], ai[RB];
#pragma unroll
    for (int r = 0; r < RB; r++) {
      const int i = min(i0 + r, n - 1);
      ki[r] = v.Kv[i];
      hi[r] = v.Hv[i];
      si[r] = v.KSv[i];
      ai[r] = v.av[i];
      if (RD == 3) {
        k2[r] = v.K2[i];
        h2[r] = v.H2[i];
        s2[r] = v.KS2[i];
      }
    }
    const int cmax = min(i0 + RB, n);
    for (int j = lane; j < cmax; j += 32) {
      const double2 kj = v.Kv[j], hj = v.Hv[j];
      const double aj = v.av[j];
      double kj2 = 0, hj2 = 0;
      if (RD == 3) {
        kj2 = v.K2[j];
        hj2 = v.H2[j];
      }
      double* Pc = c.P + plo(i0, j);
      int off = 0;
#pragma unroll
      for (int r = 0; r < RB; r++) {
        const int i = i0 + r;
        if (i < n && j <= i) {
          double d = si[r].x * kj.x;
          d = fma(-ki[r].x, hj.x, d);
          d = fma(-hi[r].x, kj.x, d);
          if (RD >= 2) {
            d = fma(si[r].y, kj.y, d);
            d = fma(-ki[r].y, hj.y, d);
            d = fma(-hi[r].y, kj.y, d);
          }
          if (RD == 3) {
            d = fma(s2[r], kj2, d);
            d = fma(-k2[r], hj2, d);
            d = fma(-h2[r], kj2, d);
          }
          const double u = fma(-ai[r], aj, 1.0);
          Pc[off] = fma(d, u, Pc[off]);
        }
        off += i + 1;  // next packed row starts i+1 further
      }
    }
  }
}

// S^-1 by partial-pivot LU on an RD x RD system held in registers (what Eigen's dynamic-size inverse() does,
// vi_ekf_meas.cpp:232; same operation order as the oracle's small_inverse), and the Mahalanobis distance.
template <int RD>
__device__ __forceinline__ double inverse_and_mahal(const double (&S)[3][3], double (&Si)[3][3], const double* res) {
  double a[RD][2 * RD];
#pragma unroll
  for (int i = 0; i < RD; i++)
#pragma unroll
    for (int j = 0; j < RD; j++) {
      a[i][j] = S[i][j];
      a[i][j + RD] = (i == j) ? 1.0 : 0.0;
    }
#pragma unroll
  for (int cc = 0; cc < RD; cc++) {
    // partial pivoting: `if (|a[i][cc]| > |a[pv][cc]|) pv = i`, rows swapped by value so every index stays static
    int pv = cc;
    double apv = fabs(a[cc][cc]);
#pragma unroll
    for (int i = cc + 1; i < RD; i++) {
      const double ai = fabs(a[i][cc]);
      if (ai > apv) {
        pv = i;
        apv = ai;
      }
    }
#pragma unroll
    for (int i = cc + 1; i < RD; i++)
      if (pv == i) {
#pragma unroll
        for (int j = 0; j < 2 * RD; j++) {
          const double t = a[cc][j];
          a[cc][j] = a[i][j];
          a[i][j] = t;
        }
      }
#pragma unroll
    for (int i = cc + 1; i < RD; i++) {
      const double f = a[i][cc] / a[cc][cc];  // the multiplier keeps the reference's rounding (it decides the pivots that follow)
#pragma unroll
      for (int j = cc; j < 2 * RD; j++) a[i][j] -= f * a[cc][j];
    }
  }
#pragma unroll
  for (int cc = RD - 1; cc >= 0; cc--) {
    const double rp = 1.0 / a[cc][cc];  // one reciprocal per pivot (1 ulp from the reference's RD divisions, far inside 1e-9)
#pragma unroll
    for (int j = 0; j < RD; j++) {
      double s = a[cc][RD + j];
#pragma unroll
      for (int k = cc + 1; k < RD; k++) s -= a[cc][k] * a[k][RD + j];
      a[cc][RD + j] = s * rp;
    }
  }
  double mahal = 0;
#pragma unroll
  for (int j = 0; j < RD; j++) {
    double s = 0;
#pragma unroll
    for (int i = 0; i < RD; i++) {
      Si[i][j] = a[i][RD + j];
      s += res[i] * a[i][RD + j];
    }
    mahal += s * res[j];
  }
  return mahal;
}

// The measurement Jacobian in registers: up to 3 rows on 6 (static) candidate columns, m of them live; h[a * 6 + k].
// One shape for every model keeps a single copy of the update code per residual dimension (instruction-cache footprint).
struct HReg {
  double h[18];
  int col[6];
  int m;
  double res[3];
  bool hnan;
};

template <int G, int RD>
__device__ int update_core(Ctx<G>& c, const KParams& p, const viekf_event& E, int n, int len, const HReg& H) {
  constexpr int MC = 6;
  int* cols = c.sc + SC_COUNT;
  UpdVec<G> v(c, 16 + 3 * p.N);
  // ---- P H^T: gather of <= MC columns of P ----
  for (int i = c.tid; i < n; i += G) {
    double s[3] = {0, 0, 0};
#pragma unroll
    for (int k = 0; k < MC; k++)
      if (k < H.m) {
        const double pv = c.P[psym(i, H.col[k])];
#pragma unroll
        for (int a = 0; a < RD; a++) s[a] = fma(pv, H.h[a * MC + k], s[a]);
      }
    v.Hv[i] = make_double2(s[0], s[1]);
    if (RD == 3) v.H2[i] = s[2];
  }
  if (c.tid == 0) cols[8] = H.hnan ? 1 : 0;
  c.sync();
  // ---- S = H P H^T + R, S^-1, Mahalanobis gate: every thread evaluates the same registers, no broadcast needed ----
  double S[3][3], Si[3][3];
  {
    double ph[MC][3];
#pragma unroll
    for (int k = 0; k < MC; k++) {
      const int ck = (k < H.m) ? H.col[k] : 0;
      const double2 t = v.Hv[ck];
      ph[k][0] = t.x;
      ph[k][1] = t.y;
      ph[k][2] = (RD == 3) ? v.H2[ck] : 0.0;
    }
#pragma unroll
    for (int a = 0; a < RD; a++)
#pragma unroll
      for (int b = 0; b < RD; b++) {
        double s = 0;
#pragma unroll
        for (int k = 0; k < MC; k++)
          if (k < H.m) s += H.h[a * MC + k] * ph[k][b];
        S[a][b] = s + E.R[a + b * RD];
      }
  }
  const double mahal = inverse_and_mahal<RD>(S, Si, H.res);
  if (mahal > 9.0) {
    if (c.tid == 0) c.sc[SC_NGATED] += 1;
    return VIEKF_MEAS_GATED;  // early return: no fix_depth (vi_ekf_meas.cpp:238)
  }
  // ---- K = P H^T S^-1, K S, state correction lambda o (K r) ----
  bool bad = false;
  for (int i = c.tid; i < n; i += G) {
    const double2 h01 = v.Hv[i];
    double ph[3] = {h01.x, h01.y, 0.0};
    if (RD == 3) ph[2] = v.H2[i];
    double k3[3] = {0, 0, 0}, ks[3] = {0, 0, 0};
#pragma unroll
    for (int b = 0; b < RD; b++) {
      double s = 0;
#pragma unroll
      for (int a = 0; a < RD; a++) s += ph[a] * Si[a][b];
      k3[b] = s;
      bad |= (s != s);
    }
    double d = 0;
#pragma unroll
    for (int a = 0; a < RD; a++) d += k3[a] * H.res[a];
#pragma unroll
    for (int a = 0; a < RD; a++) {
      double s = 0;
#pragma unroll
      for (int b = 0; b < RD; b++) s += k3[b] * S[b][a];
      ks[a] = s;
    }
    v.Kv[i] = make_double2(k3[0], k3[1]);
    v.KSv[i] = make_double2(ks[0], ks[1]);
    if (RD == 3) {
      v.K2[i] = k3[2];
      v.KS2[i] = ks[2];
    }
    const double li = c.lam[i];
    v.av[i] = p.use_partial ? 1.0 - li : 0.0;
    c.dx[i] = p.use_partial ? li * d : d;
  }
  if (bad) atomicOr((unsigned*)&cols[8], 1u);
  c.sync();
  if (!cols[8]) {
    boxplus_inplace(c, c.dx, 1.0, len);
    rank_update<G, RD>(c, v, n);
  } else if (c.tid == 0) {
    c.sc[SC_FLAGS] |= VIEKF_FLAG_NAN_GAIN;
    c.sc[SC_NNAN] += 1;
  }
  c.sync();
  for (int i = c.tid; i < len; i += G) fix_depth_one(c, p, i);
  return VIEKF_MEAS_SUCCESS;
}

// generic path: H, columns and residual were written to shared memory by thread 0
template <int G>
__device__ __forceinline__ void load_h_from_smem(Ctx<G>& c, HReg& H) {
  const double* Hc = c.sm + SM_HC;
  const double* res = c.sm + SM_RES;
  const int* cols = c.sc + SC_COUNT;
  H.m = cols[6];
  H.hnan = cols[8] != 0;
#pragma unroll
  for (int k = 0; k < 6; k++) {
    H.col[k] = cols[k];
#pragma unroll
    for (int a = 0; a < 3; a++) H.h[a * 6 + k] = Hc[a * 6 + k];
  }
#pragma unroll
  for (int a = 0; a < 3; a++) H.res[a] = res[a];
  c.sync();  // cols[8] is re-armed inside the update core
}

// The three measurement models the Monte-Carlo path lives on (h_feat vi_ekf_meas.cpp:354-367, h_acc drag model :281-296, h_alt
// :308-315) with h, H and the residual in registers.  Returns false for any other model (generic_model below).
__device__ __forceinline__ bool fast_model(const double* x, const KParams& p, int type, int rd, const double* zv, int fi, HReg& H) {
#pragma unroll
  for (int k = 0; k < 18; k++) H.h[k] = 0.0;
#pragma unroll
  for (int k = 0; k < 6; k++) H.col[k] = 0;
  H.res[0] = H.res[1] = H.res[2] = 0.0;
  H.m = 0;
  H.hnan = false;
  if (type == VIEKF_FEAT && rd == 2) {  // h_feat, vi_ekf_meas.cpp:354-367
    H.m = 2;
    H.col[0] = 16 + 3 * fi;
    H.col[1] = H.col[0] + 1;
    V3 ze, T0, T1;
    zeta_T(ldq(x + 17 + 5 * fi), ze, T0, T1);
    const double iez = 1.0 / ze.z;  // one reciprocal for the four divisions by e_z^T zeta of vi_ekf_meas.cpp:358-366 (1 ulp each)
    H.res[0] = zv[0] - (p.fl[0] * ze.x * iez + p.cc[0]);
    H.res[1] = zv[1] - (p.fl[1] * ze.y * iez + p.cc[1]);
    const double zz[3] = {ze.x, ze.y, ze.z};
    const double sk[9] = {0, -ze.z, ze.y, ze.z, 0, -ze.x, -ze.y, ze.x, 0};
    const double Tm[6] = {T0.x, T1.x, T0.y, T1.y, T0.z, T1.z};
#pragma unroll
    for (int a = 0; a < 2; a++) {
      double fb[3], fbs[3];
#pragma unroll
      for (int k = 0; k < 3; k++) fb[k] = iez * p.fl[a] * (((k == 2) ? zz[a] : 0.0) * iez - (a == k ? 1.0 : 0.0));
#pragma unroll
      for (int k = 0; k < 3; k++) fbs[k] = fb[0] * sk[k] + fb[1] * sk[3 + k] + fb[2] * sk[6 + k];
#pragma unroll
      for (int bb = 0; bb < 2; bb++) H.h[a * 6 + bb] = fbs[0] * Tm[bb] + fbs[1] * Tm[2 + bb] + fbs[2] * Tm[4 + bb];
    }
    H.hnan = (H.h[0] != H.h[0]) || (H.h[1] != H.h[1]) || (H.h[6] != H.h[6]) || (H.h[7] != H.h[7]);
    return true;
  }
  if (type == VIEKF_ACC && p.use_drag && rd == 2) {  // h_acc drag model, vi_ekf_meas.cpp:281-296
    const double mu = x[16];
    H.m = 5;
    H.col[0] = 3; H.col[1] = 4; H.col[2] = 9; H.col[3] = 10; H.col[4] = 15;
    H.h[0] = -mu; H.h[2] = 1.0; H.h[4] = -x[3];
    H.h[6 + 1] = -mu; H.h[6 + 3] = 1.0; H.h[6 + 4] = -x[4];
    H.res[0] = zv[0] - (-mu * x[3] + x[10]);
    H.res[1] = zv[1] - (-mu * x[4] + x[11]);
    H.hnan = (mu != mu) || (x[3] != x[3]) || (x[4] != x[4]);
    return true;
  }
  if (type == VIEKF_ALT && rd == 1) {  // h_alt, vi_ekf_meas.cpp:308-315
    H.m = 1;
    H.col[0] = 2;
    H.h[0] = -1.0;
    H.res[0] = zv[0] - (-x[2]);
    H.hnan = false;
    return true;
  }
  return false;
}

// Every other model (vi_ekf_meas.cpp:281-395), by one thread: H on its (sparse) columns Hc[3][6], residual res[4], cols[0..5] columns,
// cols[6] = number of live columns, cols[7] = 0, cols[8] = NaN in H.
__device__ __forceinline__ void generic_model(const double* x, const KParams& p, int type, int zdim, const double* zv, int fi, double* Hc,
                                              double* res, int* cols) {
  for (int k = 0; k < 18; k++) Hc[k] = 0.0;
  double zh[4] = {0, 0, 0, 0};
  double r4[4] = {0, 0, 0, 0};
  int m = 0;
  switch (type) {
    case VIEKF_ACC:
      if (p.use_drag) {
        const double mu = x[16];
        cols[0] = 3; cols[1] = 4; cols[2] = 9; cols[3] = 10; cols[4] = 15;
        m = 5;
        zh[0] = -mu * x[3] + x[10];
        zh[1] = -mu * x[4] + x[11];
        Hc[0] = -mu; Hc[2] = 1.0; Hc[4] = -x[3];
        Hc[6 + 1] = -mu; Hc[6 + 3] = 1.0; Hc[6 + 4] = -x[4];
      } else {
        V3 gB = rotp(ldq(x + 6), mk3(0, 0, 9.80665));
        for (int k = 0; k < 6; k++) cols[k] = 6 + k;
        m = 6;
        zh[0] = x[10] - gB.x; zh[1] = x[11] - gB.y; zh[2] = x[12] - gB.z;
        V3 ng = mk3(-gB.x, -gB.y, -gB.z);
        const double sk[9] = {0, -ng.z, ng.y, ng.z, 0, -ng.x, -ng.y, ng.x, 0};
        for (int a = 0; a < 3; a++) {
          for (int k = 0; k < 3; k++) Hc[a * 6 + k] = sk[a * 3 + k];
          Hc[a * 6 + 3 + a] = 1.0;
        }
      }
      break;
    case VIEKF_ALT:
      cols[0] = 2; m = 1; zh[0] = -x[2]; Hc[0] = -1.0;
      break;
    case VIEKF_ATT:
      for (int k = 0; k < 3; k++) { cols[k] = 6 + k; Hc[k * 6 + k] = 1.0; }
      m = 3;
      for (int k = 0; k < 4; k++) zh[k] = x[6 + k];
      break;
    case VIEKF_POS:
      for (int k = 0; k < 3; k++) { cols[k] = k; Hc[k * 6 + k] = 1.0; zh[k] = x[k]; }
      m = 3;
      break;
    case VIEKF_VEL:
      for (int k = 0; k < 3; k++) { cols[k] = 3 + k; Hc[k * 6 + k] = 1.0; zh[k] = x[3 + k]; }
      m = 3;
      break;
    case VIEKF_QZETA:
      cols[0] = 16 + 3 * fi; cols[1] = cols[0] + 1; m = 2;
      Hc[0] = 1.0; Hc[6 + 1] = 1.0;
      for (int k = 0; k < 4; k++) zh[k] = x[17 + 5 * fi + k];
      break;
    case VIEKF_FEAT: {
      cols[0] = 16 + 3 * fi; cols[1] = cols[0] + 1; m = 2;
      V3 ze, T0, T1;
      zeta_T(ldq(x + 17 + 5 * fi), ze, T0, T1);
      const double ez = ze.z;
      zh[0] = p.fl[0] * ze.x / ez + p.cc[0];
      zh[1] = p.fl[1] * ze.y / ez + p.cc[1];
      // H = (1/ez) F ((zeta e_z^T)/ez - I) [zeta]x T_zeta   (vi_ekf_meas.cpp:366)
      const double zz[3] = {ze.x, ze.y, ze.z};
      const double sk[9] = {0, -ze.z, ze.y, ze.z, 0, -ze.x, -ze.y, ze.x, 0};
      const double Tm[6] = {T0.x, T1.x, T0.y, T1.y, T0.z, T1.z};  // 3x2 row-major
      for (int a = 0; a < 2; a++) {
        double fb[3], fbs[3];
        for (int k = 0; k < 3; k++) fb[k] = (1.0 / ez) * p.fl[a] * (((k == 2) ? zz[a] : 0.0) / ez - (a == k ? 1.0 : 0.0));
        for (int k = 0; k < 3; k++) fbs[k] = fb[0] * sk[k] + fb[1] * sk[3 + k] + fb[2] * sk[6 + k];
        for (int bb = 0; bb < 2; bb++) Hc[a * 6 + bb] = fbs[0] * Tm[bb] + fbs[1] * Tm[2 + bb] + fbs[2] * Tm[4 + bb];
      }
      break;
    }
    case VIEKF_DEPTH: {
      const double rho = x[17 + 5 * fi + 4];
      cols[0] = 16 + 3 * fi + 2; m = 1;
      zh[0] = 1.0 / rho;
      Hc[0] = -1.0 / (rho * rho);
      break;
    }
    case VIEKF_INV_DEPTH:
      cols[0] = 16 + 3 * fi + 2; m = 1;
      zh[0] = x[17 + 5 * fi + 4];
      Hc[0] = 1.0;
      break;
    default:  // PIXEL_VEL: empty stub in the reference (vi_ekf_meas.cpp:388-395): zhat = 0, H = 0
      m = 0;
      break;
  }
  if (type == VIEKF_QZETA) {
    qfeat_boxminus(ldq(zv), Q4{zh[0], zh[1], zh[2], zh[3]}, r4);
  } else if (type == VIEKF_ATT) {
    V3 d = qboxminus(ldq(zv), Q4{zh[0], zh[1], zh[2], zh[3]});
    r4[0] = d.x; r4[1] = d.y; r4[2] = d.z;
  } else {
    for (int k = 0; k < zdim && k < 4; k++) r4[k] = zv[k] - zh[k];
  }
  for (int k = 0; k < 4; k++) res[k] = r4[k];
  cols[6] = m;
  cols[7] = 0;  // gate
  cols[8] = 0;  // nan
  bool hn = false;
  for (int k = 0; k < 18; k++) hn |= (Hc[k] != Hc[k]);
  if (hn) cols[8] = 1;
}

// update(meas) with add_measurement's bookkeeping in front (vi_ekf_meas.cpp:130-147,196-278). Returns meas_result_t.
// The covariance-touching steps are delegated to `core` (SmemCore below: P in shared memory), which provides
// run<RD>(E, n, len, H), init_feature(l0, l1, depth), fix_depth(len).
template <int G>
struct SmemCore {
  Ctx<G>& c;
  const KParams& p;
  template <int RD>
  __device__ __forceinline__ int run(const viekf_event& E, int n, int len, const HReg& H) {
    return update_core<G, RD>(c, p, E, n, len, H);
  }
  __device__ __forceinline__ bool init_feature(double l0, double l1, double depth) { return do_init_feature(c, p, l0, l1, depth); }
  __device__ __forceinline__ void fix_depth(int len) {
    for (int i = c.tid; i < len; i += G) fix_depth_one(c, p, i);
  }
};

template <int G, class Core>
__device__ __forceinline__ int ev_update(Ctx<G>& c, const KParams& p, const viekf_event& E, Core& core) {
  const double* zv = (E.aux == 1) ? (c.sm + SM_UB) : c.zv;
  const int type = E.type, zdim = E.zdim, rd = E.rdim;
  for (int k = 0; k < zdim; k++)
    if (zv[k] != zv[k]) return VIEKF_MEAS_NAN;
  int fi = -1;
  const bool feat_type = (type == VIEKF_FEAT || type == VIEKF_QZETA || type == VIEKF_DEPTH || type == VIEKF_INV_DEPTH);
  if (feat_type) {
    fi = (E.id >= 0) ? find_feature(c, E.id) : -1;
    if (fi < 0) {
      if (type == VIEKF_FEAT && E.id >= 0 && E.aux != 2) {  // aux == 2: update() replayed by handle_measurements, which never initialises
        core.init_feature(zv[0], zv[1], E.z[4]);
        return VIEKF_MEAS_NEW_FEATURE;
      }
      return VIEKF_MEAS_INVALID;  // the reference would index the state with -1 here (UB)
    }
  }
  const int len = c.sc[SC_LEN];
  const int n = 16 + 3 * len;
  if (!E.active) {  // logged only: no gate, no correction, but fix_depth still runs (vi_ekf_meas.cpp:230,271)
    c.sync();
    core.fix_depth(len);
    return VIEKF_MEAS_SUCCESS;
  }
  // the three models the Monte-Carlo path lives on keep h, H and the residual in registers: every thread evaluates them from
  // the same shared state, so no thread-0 section and no broadcast barrier; every other model goes through thread 0
  HReg H;
  if (!fast_model(c.x, p, type, rd, zv, fi, H)) {
    if (c.tid == 0) generic_model(c.x, p, type, zdim, zv, fi, c.sm + SM_HC, c.sm + SM_RES, c.sc + SC_COUNT);
    c.sync();
    load_h_from_smem(c, H);
  }
  // one call site per residual dimension
  if (rd == 1) return core.template run<1>(E, n, len, H);
  if (rd == 2) return core.template run<2>(E, n, len, H);
  return core.template run<3>(E, n, len, H);
}

// ------------------------------------------------------------------------------------------------------------------
// STAT
// ------------------------------------------------------------------------------------------------------------------
// e^T B^-1 e for the M x M block B of P picked by idx (M = 16: base error state, M = 6: position + attitude), by a
// right-looking Cholesky spread over the lanes of one warp (or of a W-lane segment of it, M <= W): lane i owns row i.  The subtractions hit every element in
// the same k-ascending order as a scalar left-looking factorisation.  NaN when B is not positive definite.
template <int M, int W = 32>
__device__ __forceinline__ double nees_warp(const double* P, const double* er, int lane) {
  auto idx = [](int i) { return (M == 16) ? i : (i < 3 ? i : i + 3); };
  const int li = lane < M ? lane : M - 1;
  double a[M];
#pragma unroll
  for (int j = 0; j < M; j++) a[j] = (lane < M && j <= lane) ? P[psym(idx(li), idx(j))] : 0.0;
  double r = (lane < M) ? er[idx(li)] : 0.0;
  double q = 0;
  bool bad = false;
#pragma unroll
  for (int j = 0; j < M; j++) {
    const double sjj = __shfl_sync(0xffffffffu, a[j], j, W);
    bad |= !(sjj > 0.0);
    const double ljj = sqrt(sjj);
    const double lij = (lane == j) ? ljj : a[j] / ljj;
    const double yj = __shfl_sync(0xffffffffu, r / ljj, j, W);
    q += yj * yj;
    if (lane > j) r -= lij * yj;
#pragma unroll
    for (int k = j + 1; k < M; k++) {
      const double lkj = __shfl_sync(0xffffffffu, lij, k, W);
      if (lane >= k) a[k] -= lij * lkj;
    }
  }
  return bad ? nan("") : q;
}

// NaNsInTheHouse (vi_ekf_error.cpp:6-19) for a thread group: every thread passes whether the part of the live x / P it
// scanned holds a NaN; the sticky VIEKF_FLAG_NAN_STATE is latched and returned (uniformly).
template <int G>
__device__ __forceinline__ bool latch_nan_state(const Ctx<G>& c, bool bad) {
  if (bad) atomicOr((unsigned*)&c.sc[SC_FLAGS], (unsigned)VIEKF_FLAG_NAN_STATE);
  c.sync();
  return (c.sc[SC_FLAGS] & VIEKF_FLAG_NAN_STATE) != 0;
}
// the scan of the shared-memory executor: live prefix of x and of the packed P
template <int G>
__device__ __forceinline__ bool smem_nan_scan(const Ctx<G>& c) {
  const int len = c.sc[SC_LEN];
  const int nx = 17 + 5 * len, n = 16 + 3 * len, np = n * (n + 1) / 2;
  bool bad = false;
  for (int i = c.tid; i < nx; i += G) bad |= (c.x[i] != c.x[i]);
  for (int i = c.tid; i < np; i += G) bad |= (c.P[i] != c.P[i]);
  return bad;
}

// STAT: `scan()` is the executor's thread-local NaN scan (evaluated only when the event has a valid bucket, as the oracle does).
template <int G, class Scan>
__device__ void ev_stat(Ctx<G>& c, const KParams& p, const viekf_event& E, int64_t ginst, Scan scan) {
  if (!p.stats || E.id < 0 || E.id >= p.n_buckets) return;
  c.sync();
  double* st = p.stats + (size_t)E.id * VIEKF_STAT_WIDTH;
  // a filter with a NaN in its live state or covariance is flagged, counted in [38], and kept out of every sum
  if (latch_nan_state(c, scan())) {
    if (c.tid == 0) atomicAdd(st + 38, 1.0);
    return;
  }
  double* er = c.sm + SM_ER;
  double* out = c.sm + SM_HC;  // [2] NEES16, NEES6
  if (c.tid == 0) {
    double xt[17];
    for (int i = 0; i < 8; i++) xt[i] = E.z[i];
    for (int i = 0; i < 9; i++) xt[8 + i] = E.R[i];
    if (E.noise >= 0 && p.values) {
      const double* v = p.values + c.b * p.vstride + 8 * (int64_t)E.noise;
      for (int i = 0; i < 6; i++) xt[10 + i] = v[i];
    }
    if (p.seed) {
      double bi[6];
      for (int k = 0; k < 6; k++) bi[k] = p.bias_range[k] * viekf_uniform_pm1(p.seed, (uint64_t)ginst, VIEKF_NOISE_EVENT_BIAS, k);
      Q4 qbu = ldq(p.qbu);
      V3 ba = rota(qbu, mk3(bi[0], bi[1], bi[2])), bg = rota(qbu, mk3(bi[3], bi[4], bi[5]));
      xt[10] += ba.x; xt[11] += ba.y; xt[12] += ba.z;
      xt[13] += bg.x; xt[14] += bg.y; xt[15] += bg.z;
    }
    // the global truth in the frame of the filter's current node: the inverse of get_global_pose (vi_ekf_kfr.cpp:14-22);
    // the identity map until the first keyframe reset
    {
      const FilterMeta& M = p.meta[c.b];
      const Q4 nq = ldq(M.node_q);
      const V3 pr = rotp(nq, mk3(xt[0] - M.node_t[0], xt[1] - M.node_t[1], xt[2] - M.node_t[2]));
      const Q4 qr = otimes(Q4{nq.w, -nq.x, -nq.y, -nq.z}, ldq(xt + 6));
      xt[0] = pr.x; xt[1] = pr.y; xt[2] = pr.z;
      stq(xt + 6, qr);
    }
    const double* x = c.x;
    for (int i = 0; i < 6; i++) er[i] = xt[i] - x[i];
    V3 d = qboxminus(ldq(xt + 6), ldq(x + 6));
    er[6] = d.x; er[7] = d.y; er[8] = d.z;
    for (int i = 0; i < 7; i++) er[9 + i] = xt[10 + i] - x[10 + i];
  }
  c.sync();
  // NEES by Cholesky of the 16x16 base block (warp 0) and of the 6x6 pose block (warp 1, or warp 0 again)
  if (c.tid < 32) {
    const double q16 = nees_warp<16>(c.P, er, c.tid);
    if (c.tid == 0) out[0] = q16;
  }
  if (G >= 64 ? (c.tid >= 32 && c.tid < 64) : (c.tid < 32)) {
    const double q6 = nees_warp<6>(c.P, er, c.tid & 31);
    if ((c.tid & 31) == 0) out[1] = q6;
  }
  c.sync();
  // an undefined (non-finite) NEES is left out of its sum; [37] / [39] count the defined ones
  const bool ok16 = isfinite(out[0]), ok6 = isfinite(out[1]);
  for (int k = c.tid; k < VIEKF_STAT_WIDTH; k += G) {
    double v;
    if (k == 0) v = 1.0;
    else if (k <= 16) v = er[k - 1];
    else if (k <= 32) v = er[k - 17] * er[k - 17];
    else if (k == 33) v = ok16 ? out[0] : 0.0;
    else if (k == 34) v = ok6 ? out[1] : 0.0;
    else if (k == 35) v = (double)c.sc[SC_NGATED];
    else if (k == 36) v = (double)c.sc[SC_NFIX];
    else if (k == 37) v = ok16 ? 1.0 : 0.0;
    else if (k == 39) v = ok6 ? 1.0 : 0.0;
    else continue;
    if (v != 0.0 || k <= 32) atomicAdd(st + k, v);
  }
  c.sync();
}

// Resolves component k of event e for this instance: shared z | per-instance value block, plus generated white noise and
// (PROP) the instance's constant IMU bias (include/viekf_batch.h: viekf_event.sigma, viekf_tape.seed / imu_bias_range).
__device__ __forceinline__ void resolve_values(const KParams& p, int64_t e, int64_t b, int64_t ginst, int k, double* out) {
  const viekf_event& E = p.ev[e];
  const int kind = E.kind;
  if (!(kind == VIEKF_EV_PROP || kind == VIEKF_EV_UPD || kind == VIEKF_EV_INIT)) return;
  double v = (E.noise >= 0 && p.values) ? p.values[b * p.vstride + 8 * (int64_t)E.noise + k] : E.z[k];
  if (p.seed) {
    const double sg = E.sigma[k];
    const double sgo = E.sigma[k ^ 1];
    if (sg != 0.0 || sgo != 0.0) {
      double nn[2];
      const uint32_t key = (kind == VIEKF_EV_PROP && E.aux >= 0) ? (0x80000000u | (uint32_t)E.aux) : (uint32_t)e;
      viekf_normal_pair(p.seed, (uint64_t)ginst, key, (uint32_t)(k >> 1), nn);
      v += sg * nn[k & 1];
    }
    if (kind == VIEKF_EV_PROP && k < 6 && p.bias_range[k] != 0.0)
      v += p.bias_range[k] * viekf_uniform_pm1(p.seed, (uint64_t)ginst, VIEKF_NOISE_EVENT_BIAS, k);
  }
  out[k] = v;
}

}  // namespace vk
