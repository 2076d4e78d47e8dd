// viekf_frag.cu -- tape executor for 1 <= N <= 8 (BASELINE configs[2] at N = 8) with the covariance in FP64 TENSOR-CORE FRAGMENTS.
//
// One warp owns one filter.  The 40 x 40 covariance (n = 16 + 3 len <= 40, dead rows / columns exact zeros) lives in registers as
// 5 x 5 blocks of mma.sync.m8n8k4.f64 accumulator fragments (lane (g, t) = (lane / 4, lane % 4) holds elements (g, 2t), (g, 2t + 1)
// of every block: 100 registers), and the three O(n^2) / O(n^3) operations run on the FP64 tensor-core pipe:
//   propagate   T = A P, P' = T A^T (+ G_d Qu G_d^T + Qx) block-wise, A = A_d read from the discretised rows in shared memory (17 of
//               its 25 blocks are non-zero); an accumulator block used as an operand is converted by intra-quad shuffles: element
//               (g, 4kk + t) lives in lane 4g + 2kk + t/2, register t % 2; a block of the symmetric P used as a B operand is the
//               transposed read of its mirror block (vi_ekf.cpp:302-304)
//   P H^T       every H has <= 6 non-zero columns: the owners of those columns scale their elements, a quad reduction sums them
//   rank update Joseph form as P += (1 - a a^T) o (G K^T + K G^T), G = K S / 2 - P H^T: one DMMA per block (two for a 3-row
//               residual) (vi_ekf_meas.cpp:254-265)
// Everything else -- dynamics and discretisation, the ten measurement models, S^-1 / gate, boxplus, feature bookkeeping, keyframe
// reset, statistics -- is the code of viekf_dev.cuh shared with the other executors, run by the same warp on the filter's shared-
// memory block; events that touch the covariance element-wise (init / clear / keep / keyframe reset / STAT) spill the fragments to
// the packed copy in shared memory, run the shared code and reload.  After every event the upper triangle is rebuilt from the lower
// one (fragment transposes by shuffle), so the state at an event boundary is a function of the packed lower triangle only:
// chunked = single launch and sharded = unsharded stay bit-identical.  Measured basis: tools/microbench/frag_n40_proto.cu (92 % of
// the FP64 peak on the covariance algebra).
#include "viekf_dev.cuh"

namespace vk {


namespace frg {
constexpr int NB = 5;  // 8 x 8 blocks per side: n <= 40

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// element (g, 4kk + t) of an accumulator block (c0, c1), for the lane (g, t): from lane 4g + 2kk + t/2, register t % 2
__device__ __forceinline__ double frag_elem(double c0, double c1, int src, bool odd) {
  const double x0 = __shfl_sync(0xffffffffu, c0, src), x1 = __shfl_sync(0xffffffffu, c1, src);
  return odd ? x1 : x0;
}
// does block (I, K) of A_d = [[A_bb, 0], [A_fb, blkdiag(A_ff)]] hold a non-zero?  (features straddle the 8-row block boundaries)
__host__ __device__ constexpr bool nzA(int I, int K) { return K < 2 || (I >= 2 && K >= 2 && (I - K <= 1 && K - I <= 1)); }

// A_d(r, cc) from the discretised rows of prop_discretise (Abd: 16 x 16, Afd: feature rows of 16 base columns + the own 3 x 3 block)
__device__ __forceinline__ double a_elem(const Ctx<32>& c, int r, int cc, int n) {
  if (r >= n || cc >= n) return 0.0;
  if (r < 16) return cc < 16 ? c.Abd[r * 16 + cc] : 0.0;
  const int fr = r - 16;
  if (cc < 16) return c.Afd[fr * 19 + cc];
  const int i = fr / 3, k = cc - 16 - 3 * i;
  return (k >= 0 && k < 3) ? c.Afd[fr * 19 + 16 + k] : 0.0;
}

// fragments <-> the packed lower triangle in shared memory (live prefix n only, as the load / store of the shared-memory executor)
__device__ __forceinline__ void spill(const Ctx<32>& c, int n, const double (&c0)[NB][NB], const double (&c1)[NB][NB]) {
  const int g = c.tid >> 2, t = c.tid & 3;
#pragma unroll
  for (int I = 0; I < NB; I++)
#pragma unroll
    for (int J = 0; J <= I; J++) {
      const int r = 8 * I + g, cc = 8 * J + 2 * t;
      if (r < n && cc <= r) c.P[plo(r, cc)] = c0[I][J];
      if (r < n && cc + 1 <= r) c.P[plo(r, cc + 1)] = c1[I][J];
    }
  __syncwarp();
}
__device__ __forceinline__ void reload(const Ctx<32>& c, int n, double (&c0)[NB][NB], double (&c1)[NB][NB]) {
  const int g = c.tid >> 2, t = c.tid & 3;
  __syncwarp();
#pragma unroll
  for (int I = 0; I < NB; I++)
#pragma unroll
    for (int J = 0; J < NB; J++) {
      const int r = 8 * I + g, cc = 8 * J + 2 * t;
      c0[I][J] = (r < n && cc < n) ? c.P[psym(r, cc)] : 0.0;
      c1[I][J] = (r < n && cc + 1 < n) ? c.P[psym(r, cc + 1)] : 0.0;
    }
}
// upper triangle := transpose of the lower one: destination element (g, 2t + e) of block (J, I) is element (2t + e, g) of block
// (I, J), in lane 4 (2t + e) + g / 2, register g % 2
template <bool FULL>
__device__ __forceinline__ void symmetrise(int lane, int nbl, double (&c0)[NB][NB], double (&c1)[NB][NB]) {
  const int g = lane >> 2, t = lane & 3;
  const int sa = 8 * t + (g >> 1), sb = sa + 4;
  const bool odd = g & 1;
#pragma unroll
  for (int I = 0; I < NB; I++)
#pragma unroll
    for (int J = 0; J <= I; J++) {
      if (!FULL && I >= nbl) continue;  // dead blocks are zero on both sides (uniform over the warp)
      const double x00 = __shfl_sync(0xffffffffu, c0[I][J], sa), x01 = __shfl_sync(0xffffffffu, c1[I][J], sa);
      const double x10 = __shfl_sync(0xffffffffu, c0[I][J], sb), x11 = __shfl_sync(0xffffffffu, c1[I][J], sb);
      const double t0 = odd ? x01 : x00, t1 = odd ? x11 : x10;
      if (I != J) {
        c0[J][I] = t0;
        c1[J][I] = t1;
      } else {  // inside a diagonal block the elements above the diagonal take the value of their mirror
        if (2 * t > g) c0[I][I] = t0;
        if (2 * t + 1 > g) c1[I][I] = t1;
      }
    }
}

// fix_depth (vi_ekf_helper.cpp:128-156): every lane walks the features (uniform), the owner of a diagonal element applies the
// covariance part; lane 0 keeps the flags / counters
__device__ __forceinline__ void fix_depth(const Ctx<32>& c, const KParams& p, int len, double (&c0)[NB][NB], double (&c1)[NB][NB]) {
  const int g = c.tid >> 2, t = c.tid & 3;
  __syncwarp();
  // lane i looks at feature i; nothing to fix is the common case and costs one ballot
  bool need = false;
  if (c.tid < len) {
    const double r = c.x[17 + 5 * c.tid + 4];
    need = (r != r) || (r < 0.0) || (r > 1e2);
  }
  unsigned todo = __ballot_sync(0xffffffffu, need);
  while (todo) {
    const int i = __ffs(todo) - 1;
    todo &= todo - 1;
    double* rho = c.x + 17 + 5 * i + 4;
    double r = *rho;
    const double reset = p.depth_reset;
    uint32_t fl = 0;
    int nfix = 0, kind = 0;
    double add = 0.0;
    if (r != r) {
      r = reset;
      fl |= VIEKF_FLAG_DEPTH_NAN_FIXED;
      nfix++;
    }
    if (r < 0.0) {
      const double err = reset - r;
      add = err * err;
      kind = 1;
      r = reset;
      fl |= VIEKF_FLAG_DEPTH_NEG_FIXED;
      nfix++;
    } else if (r > 1e2) {
      kind = 2;
      r = reset;
      fl |= VIEKF_FLAG_DEPTH_BIG_FIXED;
      nfix++;
    }
    __syncwarp();  // every lane has read rho
    if (kind) {
      const int d = 16 + 3 * i + 2, I = d >> 3, dg = d & 7;
      if (g == dg && t == (dg >> 1)) {
#pragma unroll
        for (int II = 2; II < NB; II++)
          if (II == I) {
            double& e = (dg & 1) ? c1[II][II] : c0[II][II];
            e = (kind == 1) ? e + add : p.P0f[2];
          }
      }
    }
    if (c.tid == 0) {
      *rho = r;
      c.sc[SC_FLAGS] |= (int)fl;
      c.sc[SC_NFIX] += nfix;
    }
    __syncwarp();
  }
}

}  // namespace frg
using namespace frg;

// resolve_values of viekf_dev.cuh with the instance's constant IMU bias taken from a per-slice cache (sm[SM_BIAS .. +5])
enum { SM_BIAS = 62 };
__device__ __forceinline__ void frag_resolve(const KParams& p, int64_t e, int64_t b, int64_t ginst, int k, const double* bias, double* out) {
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
    if (kind == VIEKF_EV_PROP && k < 6 && p.bias_range[k] != 0.0) v += bias[k];
  }
  out[k] = v;
}

// ------------------------------------------------------------------------------------------------------------------
// PROP: propagate_state (vi_ekf.cpp:262-318); dynamics, discretisation and boxplus are the shared code
// ------------------------------------------------------------------------------------------------------------------
// x <- x [+] d s (vi_ekf_helper.cpp:88-98) on one warp: the 13 additive base components one per lane, then one quaternion item per lane
// (feature i on lane i, the attitude on lane len), which share the exp / otimes code
__device__ __forceinline__ void frag_boxplus(double* __restrict__ x, const double* __restrict__ d, double s, int len, int lane) {
  if (lane < 13) {
    const int xi = lane < 6 ? lane : lane + 4, di = lane < 6 ? lane : lane + 3;
    x[xi] = fma(d[di], s, x[xi]);
  }
  if (lane <= len) {
    const bool isb = (lane == len);
    double* xq = isb ? x + 6 : x + 17 + 5 * lane;
    const Q4 q = ldq(xq);
    V3 dv;
    if (isb) {
      dv = mk3(d[6] * s, d[7] * s, d[8] * s);
    } else {
      const double* df = d + 16 + 3 * lane;
      const double d0 = df[0] * s, d1 = df[1] * s;
      const V3 T0 = rota_ex(q), T1 = rota_ey(q);
      dv = mk3(T0.x * d0 + T1.x * d1, T0.y * d0 + T1.y * d1, T0.z * d0 + T1.z * d1);
      xq[4] = fma(df[2], s, xq[4]);
    }
    const Q4 e = qexp(dv);
    const Q4 a = isb ? q : e, b = isb ? e : q;  // base: q (x) exp(v); feature: exp(T_zeta d) (x) q
    stq(xq, otimes(a, b));
  }
}

// prop_discretise of viekf_dev.cuh with the rows dealt for one warp: the 3 len feature rows in one pass (one per lane), the base rows in
// a second (9 non-trivial + 7 identity rows); the shared version's stride-32 deal runs the feature-row code twice at n = 40
__device__ __forceinline__ bool frag_discretise(Ctx<32>& c, const KParams& p, const viekf_event& E, double& dt_out, int& len_out) {
  if (c.tid == 0) {
    const double* zz = c.zv;
    Q4 qbu = ldq(p.qbu);
    V3 a = rota(qbu, mk3(zz[0], zz[1], zz[2])), g = rota(qbu, mk3(zz[3], zz[4], zz[5]));
    double* ul = c.sm + SM_UB;
    ul[0] = a.x; ul[1] = a.y; ul[2] = a.z; ul[3] = g.x; ul[4] = g.y; ul[5] = g.z;
  }
  const int started = c.sc[SC_STARTED];
  const double tcur = *c.tcur;
  const double dt = E.t - tcur;
  if (!started) {
    __syncwarp();
    if (c.tid == 0) {
      c.sc[SC_STARTED] = 1;
      *c.tcur = E.t;
    }
    __syncwarp();
    return false;
  }
  if (fabs(dt) < 1e-6) {
    __syncwarp();
    return false;
  }
  if (dt < 0) {
    if (c.tid == 0) c.sc[SC_FLAGS] |= VIEKF_FLAG_BACKWARD_DT;
    __syncwarp();
    return false;
  }
  const int len = c.sc[SC_LEN];
  const int n = 16 + 3 * len;
  const double* zv = c.zv;
  double ub[6];
  {
    Q4 qbu = ldq(p.qbu);
    V3 a = rota(qbu, mk3(zv[0], zv[1], zv[2])), g = rota(qbu, mk3(zv[3], zv[4], zv[5]));
    ub[0] = a.x; ub[1] = a.y; ub[2] = a.z; ub[3] = g.x; ub[4] = g.y; ub[5] = g.z;
  }
  base_dynamics(c, p, ub);
  for (int i = c.tid; i < len; i += 32) feature_dynamics(c, p, ub, i);
  __syncwarp();
  discretise_rows(c, dt, n, n, 16 + c.tid, 1 << 20);             // feature rows, one per lane (3 len <= 24)
  discretise_rows(c, dt, n, n, c.tid < 16 ? c.tid : n, 1 << 20);  // base rows
  __syncwarp();
  dt_out = dt;
  len_out = len;
  return true;
}

// FULL: no live-block guards (N >= 6: nearly every block is live and the branches around the DMMAs cost 9 % at N = 8)
template <bool FULL>
__device__ __forceinline__ void frag_propagate(Ctx<32>& c, const KParams& p, const viekf_event& E, double (&c0)[NB][NB], double (&c1)[NB][NB]) {
  double dt;
  int len;
  if (!frag_discretise(c, p, E, dt, len)) return;
  const int n = 16 + 3 * len;
  const int nbl = (n + 7) >> 3;  // live 8 x 8 blocks per side
  const int lane = c.tid, g = lane >> 2, t = lane & 3;
  // dense copy of blkdiag(A_ff) (3N x 3N, row stride 3N) so that every operand element of A is one load off a per-lane row base:
  // in the packed-P area (stale while the fragments hold P) at N = 8, in the AbT | T arrays (unused by this executor) below
  const int S = 3 * p.N;
  double* Aff = (9 * p.N * p.N <= 512) ? c.AbT : c.P;
  for (int i = lane; i < S * S; i += 32) Aff[i] = 0.0;
  __syncwarp();
  if (lane < len) {
#pragma unroll
    for (int a = 0; a < 3; a++)
#pragma unroll
      for (int k = 0; k < 3; k++) Aff[(3 * lane + a) * S + 3 * lane + k] = c.Afd[(3 * lane + a) * 19 + 16 + k];
  }
  __syncwarp();
  // The inner index of an 8 x 8 x 8 block product is summed in the order (0 2 4 6 | 1 3 5 7): the lane (g, t) then contributes columns
  // 2t and 2t + 1 of an accumulator block -- exactly the two elements it holds -- so an accumulator block is an A operand (and, through
  // the symmetry of P, a B operand) WITHOUT any shuffle; the other operand, read from shared memory, is indexed to match:
  // A(8I + g, 8K + 2t + kk): rows < 16 from Abd, feature rows from Afd (base columns) / Aff (feature columns); dead rows are zero
  auto a_op = [&](int I, int K, int kk) -> double {
    const int col = 8 * K + 2 * t + kk;
    if (I < 2) return c.Abd[(8 * I + g) * 16 + col];  // K < 2 here
    const int fr = 8 * (I - 2) + g;
    if (fr >= 3 * len) return 0.0;
    if (K < 2) return c.Afd[fr * 19 + col];
    const int fc = col - 16;
    return fc < S ? Aff[fr * S + fc] : 0.0;
  };
  // ---- T = A P: B operand P(K,J)[2t + kk][g] = P(J,K)[g][2t + kk], the lane's own element of the mirror block ----
  double t0[NB][NB], t1[NB][NB];
#pragma unroll
  for (int I = 0; I < NB; I++)
#pragma unroll
    for (int J = 0; J < NB; J++) t0[I][J] = t1[I][J] = 0.0;
#pragma unroll
  for (int K = 0; K < NB; K++)
#pragma unroll
    for (int kk = 0; kk < 2; kk++)
#pragma unroll
      for (int I = 0; I < NB; I++)
        if (nzA(I, K) && (FULL || (I < nbl && K < nbl))) {  // blocks behind the live prefix are zero (uniform over the warp)
          const double a = a_op(I, K, kk);
#pragma unroll
          for (int J = 0; J < NB; J++)
            if (FULL || J < nbl) dmma(t0[I][J], t1[I][J], a, kk ? c1[J][K] : c0[J][K]);
        }
  // ---- P' = T A^T + (G_d Qu) G_d^T (overwrites P): A operand T(I,K)[g][2t + kk], the lane's own element ----
#pragma unroll
  for (int I = 0; I < NB; I++)
#pragma unroll
    for (int J = 0; J < NB; J++) c0[I][J] = c1[I][J] = 0.0;
#pragma unroll
  for (int K = 0; K < NB; K++)
#pragma unroll
    for (int kk = 0; kk < 2; kk++)
#pragma unroll
      for (int J = 0; J < NB; J++)
        if (nzA(J, K) && (FULL || (J < nbl && K < nbl))) {
          const double b = a_op(J, K, kk);  // A(J,K)^T[2t + kk][g]
#pragma unroll
          for (int I = J; I < NB; I++)
            if (FULL || I < nbl) dmma(c0[I][J], c1[I][J], kk ? t1[I][K] : t0[I][K], b);  // lower blocks only: symmetrise() rebuilds the upper ones
        }
#pragma unroll
  for (int kk = 0; kk < 2; kk++) {
    const int q = 4 * kk + t;
    double a[NB], b[NB];
#pragma unroll
    for (int I = 0; I < NB; I++) {
      const int r = 8 * I + g;
      const double gd = (q < 6 && r < n) ? c.Gd[r * 6 + q] : 0.0;
      b[I] = gd;
      a[I] = (q < 6) ? gd * p.Qu[q < 6 ? q : 0] : 0.0;
    }
#pragma unroll
    for (int I = 0; I < NB; I++)
#pragma unroll
      for (int J = 0; J <= I; J++)
        if (FULL || I < nbl) dmma(c0[I][J], c1[I][J], a[I], b[J]);
  }
  // process noise on the live diagonal (vi_ekf.cpp:304)
#pragma unroll
  for (int I = 0; I < NB; I++) {
    const int r = 8 * I + g;
    if (r < n && t == (g >> 1)) {
      const double qx = (r < 16) ? p.Qx[r] : p.Qxf[(r - 16) % 3];
      if (g & 1) c1[I][I] += qx; else c0[I][I] += qx;
    }
  }
  // ---- state: x <- x [+] xdot dt, time, fix_depth ----
  frag_boxplus(c.x, c.dx, dt, len, c.tid);
  if (c.tid == 0) *c.tcur = E.t;
  fix_depth(c, p, len, c0, c1);
}

// ------------------------------------------------------------------------------------------------------------------
// UPD: the covariance-touching part of update() behind ev_update of viekf_dev.cuh (vi_ekf_meas.cpp:226-278)
// ------------------------------------------------------------------------------------------------------------------
template <bool FULL>
struct FragCore {
  Ctx<32>& c;
  const KParams& p;
  double (&c0)[NB][NB];
  double (&c1)[NB][NB];

  __device__ __forceinline__ bool init_feature(double l0, double l1, double depth) {
    const int n = 16 + 3 * c.sc[SC_LEN];
    spill(c, n, c0, c1);
    const bool ok = do_init_feature(c, p, l0, l1, depth);
    reload(c, 16 + 3 * c.sc[SC_LEN], c0, c1);
    return ok;
  }
  __device__ __forceinline__ void fix_depth(int len) { frg::fix_depth(c, p, len, c0, c1); }

  template <int RD>
  __device__ __forceinline__ int run(const viekf_event& E, int n, int len, const HReg& H) {
    constexpr int MC = 6;
    const int lane = c.tid, g = lane >> 2, t = lane & 3;
    int* cols = c.sc + SC_COUNT;
    const int nmax = 16 + 3 * p.N;  // rows of the update workspace (the fragment grid covers 40)
    const int nbl = (n + 7) >> 3;   // live 8 x 8 blocks per side
    UpdVec<32> v(c, nmax);
    // ---- P H^T: the owners of the <= 6 non-zero columns scale their elements, the quad sums ----
    double acc[NB][3];
#pragma unroll
    for (int I = 0; I < NB; I++) acc[I][0] = acc[I][1] = acc[I][2] = 0.0;
#pragma unroll
    for (int k = 0; k < MC; k++)
      if (k < H.m) {
        const int col = H.col[k], J = col >> 3, tt = (col & 7) >> 1;
        const bool e = col & 1;
        double pv[NB];
#define VIEKF_FRAG_PICK(JJ)                                      \
  _Pragma("unroll") for (int I = 0; I < NB; I++) pv[I] = e ? c1[I][JJ] : c0[I][JJ];
        switch (J) {  // uniform over the warp
          case 0: VIEKF_FRAG_PICK(0) break;
          case 1: VIEKF_FRAG_PICK(1) break;
          case 2: VIEKF_FRAG_PICK(2) break;
          case 3: VIEKF_FRAG_PICK(3) break;
          default: VIEKF_FRAG_PICK(4) break;
        }
#undef VIEKF_FRAG_PICK
        if (t != tt) {
#pragma unroll
          for (int I = 0; I < NB; I++) pv[I] = 0.0;
        }
#pragma unroll
        for (int I = 0; I < NB; I++)
#pragma unroll
          for (int a = 0; a < RD; a++) acc[I][a] = fma(pv[I], H.h[a * MC + k], acc[I][a]);
      }
#pragma unroll
    for (int I = 0; I < NB; I++) {
#pragma unroll
      for (int a = 0; a < RD; a++) {
        acc[I][a] += __shfl_xor_sync(0xffffffffu, acc[I][a], 1);
        acc[I][a] += __shfl_xor_sync(0xffffffffu, acc[I][a], 2);
      }
      const int r = 8 * I + g;
      if (t == 0 && r < n) {
        v.Hv[r] = make_double2(acc[I][0], acc[I][1]);
        if (RD == 3) v.H2[r] = acc[I][2];
      }
    }
    if (lane == 0) cols[8] = H.hnan ? 1 : 0;
    __syncwarp();
    // ---- S = H P H^T + R, S^-1, Mahalanobis gate (the code of update_core) ----
    double S[3][3], Si[3][3];
    {
      double ph[MC][3];
#pragma unroll
      for (int k = 0; k < MC; k++) {
        const int ck = (k < H.m) ? H.col[k] : 0;
        const double2 tt2 = v.Hv[ck];
        ph[k][0] = tt2.x;
        ph[k][1] = tt2.y;
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
      if (lane == 0) c.sc[SC_NGATED] += 1;
      return VIEKF_MEAS_GATED;  // early return: no fix_depth (vi_ekf_meas.cpp:238)
    }
    // ---- K = P H^T S^-1, G = K S / 2 - P H^T, lambda o (K r) ----
    bool bad = false;
    for (int i = lane; i < 8 * NB; i += 32) {
      double k3[3] = {0, 0, 0}, g3[3] = {0, 0, 0};
      double d = 0, av = 0;
      if (i < n) {
        const double2 h01 = v.Hv[i];
        double ph[3] = {h01.x, h01.y, 0.0};
        if (RD == 3) ph[2] = v.H2[i];
#pragma unroll
        for (int b = 0; b < RD; b++) {
          double s = 0;
#pragma unroll
          for (int a = 0; a < RD; a++) s += ph[a] * Si[a][b];
          k3[b] = s;
          bad |= (s != s);
        }
#pragma unroll
        for (int a = 0; a < RD; a++) d += k3[a] * H.res[a];
#pragma unroll
        for (int a = 0; a < RD; a++) {
          double s = 0;
#pragma unroll
          for (int b = 0; b < RD; b++) s += k3[b] * S[b][a];
          g3[a] = 0.5 * s - ph[a];
        }
        const double li = c.lam[i];
        av = p.use_partial ? 1.0 - li : 0.0;
        c.dx[i] = p.use_partial ? li * d : d;
      }
      // rows behind the live prefix are written as zeros: the fragment rows there are zero and stay zero
      if (i < nmax) {
        v.Kv[i] = make_double2(k3[0], k3[1]);
        v.KSv[i] = make_double2(g3[0], g3[1]);
        if (RD == 3) {
          v.K2[i] = k3[2];
          v.KS2[i] = g3[2];
        }
        v.av[i] = av;
      }
    }
    if (__any_sync(0xffffffffu, bad)) cols[8] = 1;
    __syncwarp();
    if (!cols[8]) {
      frag_boxplus(c.x, c.dx, 1.0, len, c.tid);
      // ---- P += (1 - a a^T) o (G K^T + K G^T): A operand rows (g0 g1 k0 k1), B operand columns (k0 k1 g0 g1) ----
      double a[NB], b[NB], a2[NB], b2[NB];
      // element t of the rows (g0 g1 k0 k1) / (k0 k1 g0 g1): the vectors are the interleaved pairs Kv / KSv of the update workspace
      const double* wA = c.upd + ((t < 2) ? 4 * nmax : 0) + (t & 1);
      const double* wB = c.upd + ((t < 2) ? 0 : 4 * nmax) + (t & 1);
#pragma unroll
      for (int I = 0; I < NB; I++) {
        const int r = 8 * I + g;
        const bool in = r < nmax;
        a[I] = in ? wA[2 * r] : 0.0;
        b[I] = in ? wB[2 * r] : 0.0;
        if (RD == 3) {
          const double g2 = in ? v.KS2[r] : 0.0, k2 = in ? v.K2[r] : 0.0;
          a2[I] = t == 0 ? g2 : t == 2 ? k2 : 0.0;
          b2[I] = t == 0 ? k2 : t == 2 ? g2 : 0.0;
        }
      }
      if (p.use_partial) {
#pragma unroll
        for (int I = 0; I < NB; I++) {
          const double ai = (8 * I + g < nmax) ? v.av[8 * I + g] : 0.0;
#pragma unroll
          for (int J = 0; J <= I; J++) {
            if (!FULL && I >= nbl) continue;
            double d0 = 0.0, d1 = 0.0;
            dmma(d0, d1, a[I], b[J]);
            if (RD == 3) dmma(d0, d1, a2[I], b2[J]);
            const int cc = 8 * J + 2 * t;
            const double aj0 = (cc < nmax) ? v.av[cc] : 0.0, aj1 = (cc + 1 < nmax) ? v.av[cc + 1] : 0.0;
            c0[I][J] = fma(d0, fma(-ai, aj0, 1.0), c0[I][J]);
            c1[I][J] = fma(d1, fma(-ai, aj1, 1.0), c1[I][J]);
          }
        }
      } else {
#pragma unroll
        for (int I = 0; I < NB; I++)
#pragma unroll
          for (int J = 0; J <= I; J++) {
            if (!FULL && I >= nbl) continue;
            dmma(c0[I][J], c1[I][J], a[I], b[J]);
            if (RD == 3) dmma(c0[I][J], c1[I][J], a2[I], b2[J]);
          }
      }
    } else if (lane == 0) {
      c.sc[SC_FLAGS] |= VIEKF_FLAG_NAN_GAIN;
      c.sc[SC_NNAN] += 1;
    }
    frg::fix_depth(c, p, len, c0, c1);
    return VIEKF_MEAS_SUCCESS;
  }
};

// ------------------------------------------------------------------------------------------------------------------
// the persistent tape executor (skeleton of viekf_tape_kernel<32>: same shared-memory block per filter, same load / store)
// ------------------------------------------------------------------------------------------------------------------
template <int WPC, bool FULL>
__global__ void __launch_bounds__(32 * WPC, 8 / WPC) viekf_frag_kernel(const __grid_constant__ KParams p) {
  extern __shared__ __align__(16) double smem[];
  constexpr int G = 32, GPC = WPC;
  const Layout L = make_layout(p.N);
  Ctx<32> c;
  c.gi = threadIdx.x / G;
  c.tid = threadIdx.x % G;
  double* base = smem + (size_t)c.gi * L.total;
  c.x = base + L.x; c.P = base + L.P; c.lam = base + L.lam; c.dx = base + L.dx; c.Abd = base + L.Abd; c.AbT = base + L.AbT; c.T = base + L.T;
  c.Ac = base + L.Ac; c.Gb = base + L.Gb; c.Afd = base + L.Afd; c.Gd = base + L.Gd; c.Y = base + L.Y; c.upd = base + L.upd;
  c.sm = base + L.small;
  c.ids = (int*)(base + L.ids);
  c.rm = c.ids + p.N;
  c.sc = c.rm + p.N;
  c.tcur = c.sm + SM_T;
  double* evb = smem + (size_t)GPC * L.total + c.gi * 96;
  static_assert(sizeof(viekf_event) == 31 * sizeof(double), "event staging assumes 31 doubles per record");
  const int nmax = 16 + 3 * p.N;
  for (int i = c.tid; i < nmax; i += G) c.lam[i] = (i < 16) ? p.lam[i] : p.lamf[(i - 16) % 3];
  for (int64_t b0 = (int64_t)blockIdx.x * GPC; b0 < p.B; b0 += (int64_t)gridDim.x * GPC) {
    const int64_t b = b0 + c.gi;
    if (b >= p.B) {  // ragged last round: keep the barrier count of the warps that do have a filter
      for (int64_t e = 0; e < p.ev_count; e++) __syncthreads();
      continue;
    }
    c.b = b;
    const int64_t ginst = p.inst_off + b;
    FilterMeta& M = p.meta[b];
    // ---- load the filter: meta, ids, live part of x and of the packed P; P then moves into the fragments ----
    __syncwarp();
    const int len0 = M.len;
    if (c.tid == 0) {
      c.sc[SC_LEN] = len0; c.sc[SC_NEXT] = M.next_id; c.sc[SC_STARTED] = M.started; c.sc[SC_FLAGS] = (int)M.flags;
      c.sc[SC_NGATED] = (int)M.n_gated; c.sc[SC_NFIX] = (int)M.n_fix; c.sc[SC_NNAN] = (int)M.n_nan_gain; c.sc[SC_NKF] = M.n_kf;
      *c.tcur = M.t;
      for (int k = 0; k < 6; k++) c.sm[SM_UB + k] = M.ub[k];
    }
    for (int i = c.tid; i < len0; i += G) c.ids[i] = p.ids[b * p.N + i];
    {
      const double* gx = p.x + b * p.xs;
      const int nx = 17 + 5 * len0;
      for (int i = c.tid; i < nx; i += G) c.x[i] = gx[i];
      const double2* gP = (const double2*)(p.P + b * p.ps);
      double2* sP = (double2*)c.P;
      const int n0 = 16 + 3 * len0;
      const int np2 = (n0 * (n0 + 1) / 2 + 1) >> 1;
      for (int i = c.tid; i < np2; i += G) sP[i] = gP[i];
    }
    if (c.tid < 6)
      c.sm[SM_BIAS + c.tid] = (p.seed && p.bias_range[c.tid] != 0.0) ? p.bias_range[c.tid] * viekf_uniform_pm1(p.seed, (uint64_t)ginst, VIEKF_NOISE_EVENT_BIAS, c.tid) : 0.0;
    __syncwarp();
    // the event records are staged three deep in shared memory (this event, the next one for the value resolution, one in flight)
    if (c.tid < 31) {
      if (p.ev_count > 0) evb[c.tid] = ((const double*)(p.ev + p.ev_first))[c.tid];
      if (p.ev_count > 1) evb[32 + c.tid] = ((const double*)(p.ev + p.ev_first + 1))[c.tid];
    }
    __syncwarp();
    // per-instance values are resolved four events at a time by the whole warp (8 lanes per event) into a ring of eight blocks
    // (the tail of the T array, which this executor does not use otherwise): one ~300-instruction Philox / Box-Muller pass per four events
    double* zring = c.T + 192;
    {
      const int64_t r = c.tid >> 3;
      if (r < p.ev_count) frag_resolve(p, p.ev_first + r, b, ginst, c.tid & 7, c.sm + SM_BIAS, zring + 8 * (int)(r & 7));
    }
    double c0[NB][NB], c1[NB][NB];
    reload(c, 16 + 3 * len0, c0, c1);
    __syncwarp();
    // ---- run the tape slice ----
    for (int64_t rel = 0; rel < p.ev_count; rel++) {
      const int64_t e = p.ev_first + rel;
      const viekf_event& E = *(const viekf_event*)(evb + 32 * (int)(rel % 3));
      const bool stage = rel + 2 < p.ev_count && c.tid < 31;
      double ev_next = 0.0;
      if (stage) ev_next = ((const double*)(p.ev + e + 2))[c.tid];
      const int kind = E.kind;
      c.zv = zring + 8 * (int)(rel & 7);
      if ((rel & 3) == 0) {
        const int64_t r = rel + 4 + (c.tid >> 3);
        if (r < p.ev_count) frag_resolve(p, p.ev_first + r, b, ginst, c.tid & 7, c.sm + SM_BIAS, zring + 8 * (int)(r & 7));
      }
      int result = VIEKF_MEAS_NONE;
      if (kind == VIEKF_EV_PROP) {
        frag_propagate<FULL>(c, p, E, c0, c1);
        symmetrise<FULL>(c.tid, (16 + 3 * c.sc[SC_LEN] + 7) >> 3, c0, c1);
      } else if (kind == VIEKF_EV_UPD) {
        FragCore<FULL> core{c, p, c0, c1};
        result = ev_update(c, p, E, core);
        symmetrise<FULL>(c.tid, (16 + 3 * c.sc[SC_LEN] + 7) >> 3, c0, c1);
      } else {
        // events that touch the covariance element by element run the shared code on the packed copy
        spill(c, 16 + 3 * c.sc[SC_LEN], c0, c1);
        switch (kind) {
          case VIEKF_EV_INIT:
            result = do_init_feature(c, p, c.zv[0], c.zv[1], E.z[4]) ? 1 : 0;
            break;
          case VIEKF_EV_CLEAR:
            do_clear_feature(c, p, find_feature(c, E.id));
            break;
          case VIEKF_EV_KEEP:
            do_keep_only(c, p, p.id_pool + E.ids_off, E.n_ids);
            break;
          case VIEKF_EV_KFRESET:
            do_keyframe_reset(c, p);
            break;
          case VIEKF_EV_STAT:
            ev_stat(c, p, E, ginst, [&]() { return smem_nan_scan(c); });
            break;
          default:
            break;
        }
        reload(c, 16 + 3 * c.sc[SC_LEN], c0, c1);
      }
      if (stage) evb[32 * (int)((rel + 2) % 3) + c.tid] = ev_next;
      __syncthreads();
      if (p.results && c.tid == 0 && b >= p.res_first && b < p.res_first + p.res_count)
        p.results[(b - p.res_first) * p.n_events + e] = (uint8_t)result;
    }
    // ---- store the filter ----
    spill(c, 16 + 3 * c.sc[SC_LEN], c0, c1);
    latch_nan_state(c, smem_nan_scan(c));  // NaNsInTheHouse at the end of the slice
    const int len1 = c.sc[SC_LEN];
    if (c.tid == 0) {
      M.len = len1; M.next_id = c.sc[SC_NEXT]; M.started = c.sc[SC_STARTED]; M.flags = (uint32_t)c.sc[SC_FLAGS];
      M.n_gated = (uint32_t)c.sc[SC_NGATED]; M.n_fix = (uint32_t)c.sc[SC_NFIX]; M.n_nan_gain = (uint32_t)c.sc[SC_NNAN]; M.n_kf = c.sc[SC_NKF];
      M.t = *c.tcur;
      for (int k = 0; k < 6; k++) M.ub[k] = c.sm[SM_UB + k];
    }
    for (int i = c.tid; i < len1; i += G) p.ids[b * p.N + i] = c.ids[i];
    {
      double* gx = p.x + b * p.xs;
      const int nx = 17 + 5 * len1;
      for (int i = c.tid; i < nx; i += G) gx[i] = c.x[i];
      double2* gP = (double2*)(p.P + b * p.ps);
      const double2* sP = (const double2*)c.P;
      const int n1 = 16 + 3 * len1;
      const int np2 = (n1 * (n1 + 1) / 2 + 1) >> 1;
      for (int i = c.tid; i < np2; i += G) gP[i] = sP[i];
    }
    __syncwarp();
  }
}

// filters (= warps) per CTA, in event lock-step.  Eight warps fit an SM (255 registers): as ONE CTA they share the instruction stream
// (40 M against 36 M filter-steps/s as two CTAs of four at N = 8, 46.7 M against 43.3 M at N = 4)
static int frag_wpc(int N) { return 8; }
size_t frag_smem_bytes(int N) { return (smem_bytes_per_filter(N) + 96 * sizeof(double)) * frag_wpc(N); }

template <int WPC, bool FULL>
static cudaError_t launch_frag_wpc(const KParams& p, int sm_count, cudaStream_t stream) {
  const size_t smem = frag_smem_bytes(p.N);
  cudaError_t e = cudaFuncSetAttribute(viekf_frag_kernel<WPC, FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(viekf_frag_kernel<WPC, FULL>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return e;
  int occ = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, viekf_frag_kernel<WPC, FULL>, 32 * WPC, smem);
  if (e != cudaSuccess) return e;
  if (occ < 1) return cudaErrorInvalidConfiguration;
  const int64_t need = (p.B + WPC - 1) / WPC;
  int64_t grid = (int64_t)sm_count * occ;  // persistent: a whole number of waves of the SM count
  if (need < grid) grid = need;
  if (grid < 1) grid = 1;
  viekf_frag_kernel<WPC, FULL><<<(unsigned)grid, 32 * WPC, smem, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_frag(const KParams& p, int sm_count, cudaStream_t stream) {
  if (p.N < 1 || p.N > 8) return cudaErrorInvalidValue;
  return p.N >= 6 ? launch_frag_wpc<8, true>(p, sm_count, stream) : launch_frag_wpc<8, false>(p, sm_count, stream);
}

}  // namespace vk
