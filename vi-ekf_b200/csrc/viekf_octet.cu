// viekf_octet.cu -- tape executor for N = 0 (BASELINE configs[1]: IMU propagation + base-state updates), sm_100a, FP64.
//
// Eight lanes own one filter, four filters share a warp: the 16 x 16 covariance lives in REGISTERS (lane s of the octet holds
// the full rows s and s + 8, 32 doubles), the state and the small per-event matrices in shared memory (3.8 KB per filter, so
// 48 filters are resident per SM where the warp-per-filter mapping of viekf_kernel.cu holds 12).  Every hand-off inside an event
// is a __syncwarp(); the filters of a warp run the shared tape in lock-step by construction, and the warps of a CTA meet at one
// barrier per event so that they share the instruction stream.  Event semantics are those of viekf_kernel.cu (same reference citations):
//   PROP    VIEKF::propagate_state + dynamics + boxplus                 (vi_ekf.cpp:262-318, vi_ekf_dyn.cpp:14-94)
//   UPD     add_measurement bookkeeping + VIEKF::update + base h-models (vi_ekf_meas.cpp:130-147,196-352)
//   INIT / CLEAR / KEEP  with NUM_FEATURES = 0: refuse / no-op / keyframe bookkeeping only (vi_ekf_feat.cpp:6-142)
//   KFRESET keyframe_reset (vi_ekf_kfr.cpp:56-157),  STAT  RMSE / NEES partial sums
//
// Covariance propagation with A_d = I + M, M = A dt + A^2 dt^2 / 2 (rows 9..15 and columns 0..2 of M are zero):
//   U = P (I + M)^T       own rows, columns 0..8 change:   U[i][c] = P[i][c] + sum_k P[i][k] M[c][k]
//   P' = (I + M) U        rows 0..8 change:                P'[r][j] = U[r][j] + sum_k M[r][k] U[k][j]
// Only the lower triangle of the 9 x 9 block is evaluated and mirrored, rows >= 9 keep U, so P stays symmetric by
// construction like the packed layouts of the other executors.
#include "viekf_dev.cuh"

namespace vk {

#ifndef VIEKF_OCTET_WPC
#define VIEKF_OCTET_WPC 4  // warps per CTA
#endif
#ifndef VIEKF_OCTET_LOCKSTEP
#define VIEKF_OCTET_LOCKSTEP 1  // the warps of a CTA run the tape in event lock-step (one CTA barrier per event): they share the
                                // instruction stream, 365 M -> 405 M filter-steps/s at 65,536 x 0 (stall_no_instruction was 3.4 of 12 cycles)
#endif
#ifndef VIEKF_OCTET_MINB
#define VIEKF_OCTET_MINB 3  // CTAs per SM the register allocation must admit
#endif

namespace oct {

// shared-memory layout of one filter, in doubles
enum {
  O_X = 0,     // x[17] (+1)
  O_UB = 18,   // body-frame IMU input of the latest propagate_state call [6]
  O_ZV = 24,   // this event's resolved values [8]
  O_DX = 32,   // xdot (PROP) / lambda o K r (UPD) [16]
  O_M = 48,    // M rows 0..8, columns 3..15 at stride 14 [126]; STAT / KFRESET scratch (er[16], Nb[9], out[2]) overlays it
  O_GD = 174,  // G_d rows 0..8 [54]
  O_X2 = 228,  // lower triangle of the new 9 x 9 block, stride 10 [90]
  O_XA = 318,  // exchange area [160]: continuous A rows (9 x 14) | U columns 0..8 (16 x 10) | P H^T, K (16 x 4 each) | packed P
  O_BIAS = 478,  // the instance's constant IMU bias (viekf_tape.imu_bias_range), drawn once per slice [6]
  O_TOTAL = 484
};
enum { MS = 14, XS = 10 };

struct Flt {  // per-filter scalars, replicated in the eight lanes of the octet
  double t;
  int started, nkf, flags, ngated, nnan;
};

__device__ __forceinline__ unsigned octet_mask(int lane) { return 0xffu << (lane & 24); }

}  // namespace oct
using namespace oct;

// ------------------------------------------------------------------------------------------------------------------
// rows of P between registers and the packed global / shared layout
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void rows_from_packed(const double* Pk, int sub, double (&Pa)[16], double (&Pb)[16]) {
#pragma unroll
  for (int j = 0; j < 16; j++) {
    Pa[j] = Pk[psym(sub, j)];
    Pb[j] = Pk[plo(sub + 8, 0) + (j <= sub + 8 ? j : 0)];
  }
#pragma unroll
  for (int j = 9; j < 16; j++)
    if (j > sub + 8) Pb[j] = Pk[plo(j, sub + 8)];
}
__device__ __forceinline__ void rows_to_packed(double* Pk, int sub, const double (&Pa)[16], const double (&Pb)[16]) {
#pragma unroll
  for (int j = 0; j < 8; j++)
    if (j <= sub) Pk[plo(sub, j)] = Pa[j];
#pragma unroll
  for (int j = 0; j < 16; j++)
    if (j <= sub + 8) Pk[plo(sub + 8, j)] = Pb[j];
}

// Component k of the event's per-instance values (resolve_values of viekf_dev.cuh on the staged record, the instance's IMU bias
// from the per-slice cache): shared z | per-instance value block, plus generated white noise and (PROP) the constant bias.
__device__ __forceinline__ void oct_resolve(const KParams& p, const viekf_event& E, int64_t e, int64_t b, int64_t ginst, int k, double* s) {
  const int kind = E.kind;
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
    if (kind == VIEKF_EV_PROP && k < 6 && p.bias_range[k] != 0.0) v += s[O_BIAS + k];
  }
  s[O_ZV + k] = v;
}

// ------------------------------------------------------------------------------------------------------------------
// PROP
// ------------------------------------------------------------------------------------------------------------------
// x <- x [+] d s for the base state (vi_ekf_helper.cpp:88-98), one component per lane: lanes 0..5 POS / VEL and B_A / B_G,
// lane 6 the quaternion, lane 7 MU.  `go` masks filters that do not apply the step.
__device__ __forceinline__ void oct_boxplus(double* s, const double* d, double sc, int sub, bool go) {
  double* x = s + O_X;
  if (go) {
    if (sub < 6) {
      x[sub] += d[sub] * sc;
      x[10 + sub] += d[9 + sub] * sc;
    } else if (sub == 6) {
      stq(x + 6, otimes(ldq(x + 6), qexp(mk3(d[6] * sc, d[7] * sc, d[8] * sc))));
    } else {
      x[16] += d[15] * sc;
    }
  }
}

__device__ __forceinline__ void oct_propagate(double* s, const KParams& p, const viekf_event& E, int sub, Flt& F, double (&Pa)[16],
                                           double (&Pb)[16]) {
  const double* zv = s + O_ZV;
  // the body-frame input of the latest propagate_state call is kept for UPD events with aux == 1 (vi_ekf.cpp:265-267)
  double ub[6];
  {
    Q4 qbu = ldq(p.qbu);
    V3 a = rota(qbu, mk3(zv[0], zv[1], zv[2])), g = rota(qbu, mk3(zv[3], zv[4], zv[5]));
    ub[0] = a.x; ub[1] = a.y; ub[2] = a.z; ub[3] = g.x; ub[4] = g.y; ub[5] = g.z;
    if (sub < 6) s[O_UB + sub] = sub == 0 ? a.x : sub == 1 ? a.y : sub == 2 ? a.z : sub == 3 ? g.x : sub == 4 ? g.y : g.z;
  }
  // guards of propagate_state (vi_ekf.cpp:274-289), per filter
  const double dt = E.t - F.t;
  bool adv = false;
  if (!F.started) {
    F.started = 1;
    F.t = E.t;
  } else if (fabs(dt) < 1e-6) {
  } else if (dt < 0) {
    F.flags |= VIEKF_FLAG_BACKWARD_DT;
  } else {
    adv = true;
  }
  if (!__any_sync(0xffffffffu, adv)) {
    __syncwarp();
    return;
  }
  double* Ac = s + O_XA;
  // ---- phase A: continuous dynamics and the non-zero entries of A rows POS / VEL / ATT, columns 3..15 (vi_ekf_dyn.cpp:14-94) ----
  {
    double2* z2 = (double2*)Ac;
#pragma unroll
    for (int k = 0; k < 8; k++) z2[sub * 8 + k] = make_double2(0.0, 0.0);
  }
  __syncwarp();
  const double* x = s + O_X;
  const V3 vel = ld3(x + 3);
  if (sub == 0) {
    const Q4 q = ldq(x + 6);
    const V3 acc = mk3(ub[0] - x[10], ub[1] - x[11], ub[2] - x[12]);
    const V3 om = mk3(ub[3] - x[13], ub[4] - x[14], ub[5] - x[15]);
    const double mu = x[16];
    double R[9];
    Rmat(q, R);
    const V3 gB = rotp(q, mk3(0, 0, 9.80665));
    const V3 pd = rota(q, vel);
    const V3 oxv = cross(om, vel);
    V3 vd;
    if (p.use_drag)
      vd = mk3(gB.x - oxv.x - mu * vel.x, gB.y - oxv.y - mu * vel.y, acc.z + gB.z - oxv.z);
    else
      vd = acc + gB - oxv;
    double* dx = s + O_DX;
    dx[0] = pd.x; dx[1] = pd.y; dx[2] = pd.z;
    dx[3] = vd.x; dx[4] = vd.y; dx[5] = vd.z;
    dx[6] = om.x; dx[7] = om.y; dx[8] = om.z;
#pragma unroll
    for (int i = 9; i < 16; i++) dx[i] = 0.0;
    const double sv[9] = {0, -vel.z, vel.y, vel.z, 0, -vel.x, -vel.y, vel.x, 0};
    const double so[9] = {0, -om.z, om.y, om.z, 0, -om.x, -om.y, om.x, 0};
    const double sg[9] = {0, -gB.z, gB.y, gB.z, 0, -gB.x, -gB.y, gB.x, 0};
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
      for (int cc = 0; cc < 3; cc++) {
        Ac[r * MS + cc] = R[cc * 3 + r];  // (POS,VEL) = R^T
        double t = 0;
#pragma unroll
        for (int k = 0; k < 3; k++) t += (-R[k * 3 + r]) * sv[k * 3 + cc];
        Ac[r * MS + 3 + cc] = t;  // (POS,ATT) = -R^T [v]x
        if (r != cc) {
          Ac[(3 + r) * MS + cc] = -so[r * 3 + cc];
          Ac[(3 + r) * MS + 3 + cc] = sg[r * 3 + cc];
          Ac[(3 + r) * MS + 9 + cc] = -sv[r * 3 + cc];
          Ac[(6 + r) * MS + 3 + cc] = -so[r * 3 + cc];
        }
      }
#pragma unroll
    for (int r = 0; r < 3; r++) Ac[(6 + r) * MS + 9 + r] = -1.0;
    if (p.use_drag) {
      Ac[3 * MS + 0] = -mu;
      Ac[4 * MS + 1] = -mu;
      Ac[5 * MS + 8] = -1.0;
      Ac[3 * MS + 12] = -vel.x;
      Ac[4 * MS + 12] = -vel.y;
    } else {
#pragma unroll
      for (int r = 0; r < 3; r++) Ac[(3 + r) * MS + 6 + r] = -1.0;
    }
  }
  __syncwarp();
  // ---- phase B: row r of M = A dt + A^2 dt^2 / 2 and of G_d = (I + A dt / 2 + A^2 dt^2 / 6) G dt (vi_ekf.cpp:302-303) ----
  {
    const double h2 = dt * dt / 2.0, h6 = dt * dt / 6.0, hd2 = dt / 2.0;
    {
      const int r = sub;
      double ar[13];
#pragma unroll
      for (int k = 0; k < 13; k++) ar[k] = Ac[r * MS + k];
      double mrow[13], mf[6];
#pragma unroll
      for (int c0 = 0; c0 < 14; c0 += 2) {
        double a0 = 0, a1 = 0;
#pragma unroll
        for (int k = 0; k < 6; k++) {
          const double2 av = *(const double2*)(Ac + (3 + k) * MS + c0);
          a0 = fma(ar[k], av.x, a0);
          a1 = fma(ar[k], av.y, a1);
        }
        mrow[c0] = ar[c0] * dt + a0 * h2;
        if (c0 < 6) mf[c0] = ((r == 3 + c0) ? 1.0 : 0.0) + ar[c0] * hd2 + a0 * h6;
        if (c0 + 1 < 13) {
          mrow[c0 + 1] = ar[c0 + 1] * dt + a1 * h2;
          if (c0 + 1 < 6) mf[c0 + 1] = ((r == 4 + c0) ? 1.0 : 0.0) + ar[c0 + 1] * hd2 + a1 * h6;
        }
      }
      double* Mr = s + O_M + r * MS;
#pragma unroll
      for (int k = 0; k < 13; k++) Mr[k] = mrow[k];
      Mr[13] = 0.0;
      double gd[6];
      if (p.use_drag) {
        gd[0] = 0.0;
        gd[1] = 0.0;
        gd[2] = -dt * mf[2];
      } else {
#pragma unroll
        for (int q = 0; q < 3; q++) gd[q] = -dt * mf[q];
      }
      const double sv[9] = {0, -vel.z, vel.y, vel.z, 0, -vel.x, -vel.y, vel.x, 0};
#pragma unroll
      for (int cc = 0; cc < 3; cc++) {
        double t = 0;
#pragma unroll
        for (int k = 0; k < 3; k++) t = fma(-mf[k], sv[k * 3 + cc], t);
        gd[3 + cc] = dt * (t - mf[3 + cc]);
      }
      double* Gr = s + O_GD + r * 6;
#pragma unroll
      for (int q = 0; q < 6; q++) Gr[q] = gd[q];
    }
    // row 8 (ATT z) is shared out by columns, two per lane; its part of I + A dt / 2 + A^2 dt^2 / 6 goes through a scratch row
    if (sub < 7) {
      const int c0 = 2 * sub;
      double a0 = 0, a1 = 0;
#pragma unroll
      for (int k = 0; k < 6; k++) {
        const double a8 = Ac[8 * MS + k];
        const double2 av = *(const double2*)(Ac + (3 + k) * MS + c0);
        a0 = fma(a8, av.x, a0);
        a1 = fma(a8, av.y, a1);
      }
      const double2 a8c = *(const double2*)(Ac + 8 * MS + c0);
      *(double2*)(s + O_M + 8 * MS + c0) = make_double2(a8c.x * dt + a0 * h2, a8c.y * dt + a1 * h2);
      if (sub < 3) *(double2*)(s + O_X2 + c0) = make_double2(a8c.x * hd2 + a0 * h6, ((c0 == 4) ? 1.0 : 0.0) + a8c.y * hd2 + a1 * h6);
    }
  }
  __syncwarp();
  if (sub == 7) {  // G_d row 8 from the scratch row (read before the barrier that precedes the first write of X2)
    const double* mf = s + O_X2;
    const double sv[9] = {0, -vel.z, vel.y, vel.z, 0, -vel.x, -vel.y, vel.x, 0};
    double* Gr = s + O_GD + 8 * 6;
    if (p.use_drag) {
      Gr[0] = 0.0;
      Gr[1] = 0.0;
      Gr[2] = -dt * mf[2];
    } else {
#pragma unroll
      for (int q = 0; q < 3; q++) Gr[q] = -dt * mf[q];
    }
#pragma unroll
    for (int cc = 0; cc < 3; cc++) {
      double t = 0;
#pragma unroll
      for (int k = 0; k < 3; k++) t = fma(-mf[k], sv[k * 3 + cc], t);
      Gr[3 + cc] = dt * (t - mf[3 + cc]);
    }
  }
  // ---- U = P (I + M)^T on the own rows, columns 0..8 ----
  double* X = s + O_XA;
  {
#pragma unroll
    for (int c = 0; c < 9; c++) {
      double mc[14];
#pragma unroll
      for (int k = 0; k < 14; k += 2) {
        const double2 v = *(const double2*)(s + O_M + c * MS + k);
        mc[k] = v.x;
        mc[k + 1] = v.y;
      }
      double a = Pa[c], b = Pb[c];
#pragma unroll
      for (int k = 0; k < 13; k++) {
        a = fma(Pa[3 + k], mc[k], a);
        b = fma(Pb[3 + k], mc[k], b);
      }
      // straight to the exchange area: the rows keep their old columns 3..8 (operands of the columns still to come)
      X[sub * XS + c] = a;
      X[(sub + 8) * XS + c] = b;
    }
  }
  __syncwarp();
  // ---- lower triangle of the new 9 x 9 block: row `sub` (columns 0..sub) and element (8, sub) per lane, (8, 8) on lane 0 ----
  double* X2 = s + O_X2;
  {
    double mr[13], m8[13], gr[6], g8[6];
#pragma unroll
    for (int k = 0; k < 13; k++) {
      mr[k] = s[O_M + sub * MS + k];
      m8[k] = s[O_M + 8 * MS + k];
    }
#pragma unroll
    for (int q = 0; q < 6; q++) {
      gr[q] = s[O_GD + sub * 6 + q] * p.Qu[q];
      g8[q] = s[O_GD + 8 * 6 + q] * p.Qu[q];
    }
    double v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) v[j] = X[sub * XS + j];
    double v8 = X[8 * XS + sub], v88 = X[8 * XS + 8];
#pragma unroll
    for (int k = 0; k < 13; k++) {
      const double* Xr = X + (3 + k) * XS;
#pragma unroll
      for (int j = 0; j < 8; j += 2) {
        const double2 u = *(const double2*)(Xr + j);
        v[j] = fma(mr[k], u.x, v[j]);
        v[j + 1] = fma(mr[k], u.y, v[j + 1]);
      }
      v8 = fma(m8[k], Xr[sub], v8);
      v88 = fma(m8[k], Xr[8], v88);
    }
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const double* Gj = s + O_GD + j * 6;
#pragma unroll
      for (int q = 0; q < 6; q++) v[j] = fma(gr[q], Gj[q], v[j]);
    }
    {
      const double* Gs = s + O_GD + sub * 6;
      const double* G8 = s + O_GD + 8 * 6;
#pragma unroll
      for (int q = 0; q < 6; q++) {
        v8 = fma(g8[q], Gs[q], v8);
        v88 = fma(g8[q], G8[q], v88);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; j++)
      if (j <= sub) X2[sub * XS + j] = v[j] + (j == sub ? p.Qx[j] : 0.0);
    X2[8 * XS + sub] = v8;
    if (sub == 0) X2[8 * XS + 8] = v88 + p.Qx[8];
  }
  __syncwarp();
  // ---- assemble the rows: mirrored 9 x 9 block, columns >= 9 of rows < 9 from U^T, rows >= 9 keep U, process noise on the diagonal ----
  if (adv) {
#pragma unroll
    for (int c = 0; c < 9; c++) Pb[c] = X[(sub + 8) * XS + c];  // (row 8 of lane 0 is replaced below)
#pragma unroll
    for (int j = 0; j < 9; j++) Pa[j] = X2[(j > sub ? j : sub) * XS + (j > sub ? sub : j)];
#pragma unroll
    for (int j = 9; j < 16; j++) Pa[j] = X[j * XS + sub];
    if (sub == 0) {
#pragma unroll
      for (int j = 0; j < 9; j++) Pb[j] = X2[8 * XS + j];
#pragma unroll
      for (int j = 9; j < 16; j++) Pb[j] = X[j * XS + 8];
    }
#pragma unroll
    for (int j = 9; j < 16; j++)
      if (j == sub + 8) Pb[j] += p.Qx[j];
    F.t = E.t;
  }
  // ---- state: x <- x [+] xdot dt ----
  oct_boxplus(s, s + O_DX, dt, sub, adv);
  __syncwarp();
}

// ------------------------------------------------------------------------------------------------------------------
// UPD
// ------------------------------------------------------------------------------------------------------------------
// column sets of the base-state measurement models (vi_ekf_meas.cpp:281-352): static, so the rows of P stay in registers
enum { CS_NONE = 0, CS_ACC_DRAG, CS_ACC_GRAV, CS_ALT, CS_ATT, CS_POS, CS_VEL };

// P H^T for one row of P: out[a] = sum_k row[col_k] H[a][k]
template <int CS>
__device__ __forceinline__ void row_pht(const double (&row)[16], const HReg& H, double (&o)[3]) {
  constexpr int m = (CS == CS_ACC_DRAG) ? 5 : (CS == CS_ACC_GRAV) ? 6 : (CS == CS_ALT) ? 1 : (CS == CS_NONE) ? 0 : 3;
#pragma unroll
  for (int a = 0; a < 3; a++) {
    double t = 0;
#pragma unroll
    for (int k = 0; k < m; k++) {
      constexpr int dcol[5] = {3, 4, 9, 10, 15};
      const int col = (CS == CS_ACC_DRAG) ? dcol[k < 5 ? k : 0] : (CS == CS_ACC_GRAV) ? 6 + k : (CS == CS_ALT) ? 2 : (CS == CS_ATT) ? 6 + k : (CS == CS_POS) ? k : 3 + k;
      t = fma(row[col], H.h[a * 6 + k], t);
    }
    o[a] = t;
  }
}

// update() behind the gate (vi_ekf_meas.cpp:226-278) for residual dimension RD.  Returns meas_result_t.
template <int RD>
__device__ __forceinline__ int oct_update_core(double* s, const KParams& p, const viekf_event& E, int sub, int lane, Flt& F, double (&Pa)[16],
                                               double (&Pb)[16], const HReg& H, int cs, bool live) {
  double* W = s + O_XA;        // P H^T, 16 rows x 4
  double* Kb = s + O_XA + 64;  // K, 16 rows x 4
  double* Gv = s + O_X2;       // G = K S / 2 - P H^T, 16 rows x 4
  double pha[3], phb[3];
  switch (cs) {
    case CS_ACC_DRAG: row_pht<CS_ACC_DRAG>(Pa, H, pha); row_pht<CS_ACC_DRAG>(Pb, H, phb); break;
    case CS_ACC_GRAV: row_pht<CS_ACC_GRAV>(Pa, H, pha); row_pht<CS_ACC_GRAV>(Pb, H, phb); break;
    case CS_ALT: row_pht<CS_ALT>(Pa, H, pha); row_pht<CS_ALT>(Pb, H, phb); break;
    case CS_ATT: row_pht<CS_ATT>(Pa, H, pha); row_pht<CS_ATT>(Pb, H, phb); break;
    case CS_POS: row_pht<CS_POS>(Pa, H, pha); row_pht<CS_POS>(Pb, H, phb); break;
    case CS_VEL: row_pht<CS_VEL>(Pa, H, pha); row_pht<CS_VEL>(Pb, H, phb); break;
    default: row_pht<CS_NONE>(Pa, H, pha); row_pht<CS_NONE>(Pb, H, phb); break;
  }
#pragma unroll
  for (int a = 0; a < 3; a++) {
    W[sub * 4 + a] = pha[a];
    W[(sub + 8) * 4 + a] = phb[a];
  }
  __syncwarp();
  // ---- S = H P H^T + R, S^-1, Mahalanobis gate: every lane of the octet evaluates the same values ----
  double S[3][3], Si[3][3];
#pragma unroll
  for (int a = 0; a < RD; a++)
#pragma unroll
    for (int b = 0; b < RD; b++) S[a][b] = E.R[a + b * RD];
#pragma unroll
  for (int k = 0; k < 6; k++)
    if (k < H.m) {
      const double* Wr = W + H.col[k] * 4;
#pragma unroll
      for (int a = 0; a < RD; a++)
#pragma unroll
        for (int b = 0; b < RD; b++) S[a][b] = fma(H.h[a * 6 + k], Wr[b], S[a][b]);
    }
  const double mahal = inverse_and_mahal<RD>(S, Si, H.res);
  const bool gated = mahal > 9.0;
  // ---- K = P H^T S^-1, K S, lambda o (K r) on the own rows ----
  double ka[3] = {0, 0, 0}, kb[3] = {0, 0, 0}, ksa[3] = {0, 0, 0}, ksb[3] = {0, 0, 0};
  bool bad = H.hnan;
  double da = 0, db = 0;
#pragma unroll
  for (int b = 0; b < RD; b++) {
    double ta = 0, tb = 0;
#pragma unroll
    for (int a = 0; a < RD; a++) {
      ta += pha[a] * Si[a][b];
      tb += phb[a] * Si[a][b];
    }
    ka[b] = ta;
    kb[b] = tb;
    bad |= (ta != ta) || (tb != tb);
    da += ta * H.res[b];
    db += tb * H.res[b];
  }
#pragma unroll
  for (int a = 0; a < RD; a++) {
    double ta = 0, tb = 0;
#pragma unroll
    for (int b = 0; b < RD; b++) {
      ta += ka[b] * S[b][a];
      tb += kb[b] * S[b][a];
    }
    ksa[a] = ta;
    ksb[a] = tb;
  }
  // lambda of the own rows (p.lam is read through the constant bank)
  const double la = p.lam[sub], lb = p.lam[sub + 8];
  const double aa = p.use_partial ? 1.0 - la : 0.0, ab = p.use_partial ? 1.0 - lb : 0.0;
  // Joseph form as a symmetric rank update: with G = K S / 2 - P H^T,
  //   (I - K H) P (I - K H)^T + K R K^T - P = K S K^T - K (P H^T)^T - (P H^T) K^T = G K^T + K G^T
  // (K S K^T is symmetric), and every element is evaluated as sum_a (g_i[a] k_j[a] + k_i[a] g_j[a]) with separately rounded
  // products, so the owners of (i, j) and (j, i) get the same bits: P stays symmetric without an exchange.
  double ga[3], gb[3];
#pragma unroll
  for (int a = 0; a < 3; a++) {
    ga[a] = 0.5 * ksa[a] - pha[a];
    gb[a] = 0.5 * ksb[a] - phb[a];
    Kb[sub * 4 + a] = ka[a];
    Kb[(sub + 8) * 4 + a] = kb[a];
    Gv[sub * 4 + a] = ga[a];
    Gv[(sub + 8) * 4 + a] = gb[a];
  }
  s[O_DX + sub] = p.use_partial ? la * da : da;
  s[O_DX + sub + 8] = p.use_partial ? lb * db : db;
  const unsigned badm = __ballot_sync(0xffffffffu, bad) & octet_mask(lane);
  __syncwarp();  // K, G and dx are published
  int result = VIEKF_MEAS_SUCCESS;
  if (!live) {
    result = VIEKF_MEAS_NAN;  // NaN in z: rejected by add_measurement (vi_ekf_meas.cpp:136-139)
  } else if (gated) {
    F.ngated += 1;
    result = VIEKF_MEAS_GATED;  // early return of the reference: nothing applied (vi_ekf_meas.cpp:238)
  } else if (badm) {
    F.flags |= VIEKF_FLAG_NAN_GAIN;
    F.nnan += 1;
  } else {
#pragma unroll
    for (int j = 0; j < 16; j++) {
      const double2 k01 = *(const double2*)(Kb + j * 4), g01 = *(const double2*)(Gv + j * 4);
      const double aj = p.use_partial ? 1.0 - p.lam[j] : 0.0;
      double d0 = __dadd_rn(__dmul_rn(ga[0], k01.x), __dmul_rn(ka[0], g01.x));
      double d1 = __dadd_rn(__dmul_rn(gb[0], k01.x), __dmul_rn(kb[0], g01.x));
      if (RD >= 2) {
        d0 = __dadd_rn(d0, __dadd_rn(__dmul_rn(ga[1], k01.y), __dmul_rn(ka[1], g01.y)));
        d1 = __dadd_rn(d1, __dadd_rn(__dmul_rn(gb[1], k01.y), __dmul_rn(kb[1], g01.y)));
      }
      if (RD == 3) {
        const double k2 = Kb[j * 4 + 2], g2 = Gv[j * 4 + 2];
        d0 = __dadd_rn(d0, __dadd_rn(__dmul_rn(ga[2], k2), __dmul_rn(ka[2], g2)));
        d1 = __dadd_rn(d1, __dadd_rn(__dmul_rn(gb[2], k2), __dmul_rn(kb[2], g2)));
      }
      Pa[j] = fma(d0, fma(-aa, aj, 1.0), Pa[j]);
      Pb[j] = fma(d1, fma(-ab, aj, 1.0), Pb[j]);
    }
  }
  oct_boxplus(s, s + O_DX, 1.0, sub, live && !gated && !badm);
  __syncwarp();
  return result;
}

static __device__ __noinline__ void oct_generic_model(double* s, const KParams& p, int type, int zdim, const double* zv) {
  generic_model(s + O_X, p, type, zdim, zv, -1, s + O_M, s + O_M + 18, (int*)(s + O_M + 24));
}
// add_measurement bookkeeping + update (vi_ekf_meas.cpp:130-147,196-278) with NUM_FEATURES = 0
__device__ __forceinline__ int oct_update(double* s, const KParams& p, const viekf_event& E, int sub, int lane, Flt& F, double (&Pa)[16],
                                          double (&Pb)[16]) {
  const double* zv = (E.aux == 1) ? (s + O_UB) : (s + O_ZV);
  const int type = E.type, zdim = E.zdim, rd = E.rdim;
  bool live = true;
  for (int k = 0; k < zdim; k++) live &= !(zv[k] != zv[k]);
  // uniform over the batch from here on: the decisions depend on the event only
  if (type == VIEKF_FEAT || type == VIEKF_QZETA || type == VIEKF_DEPTH || type == VIEKF_INV_DEPTH) {
    // no feature can be tracked: a FEAT measurement asks init_feature, which refuses (vi_ekf_feat.cpp:9), and reports NEW_FEATURE
    if (!live) return VIEKF_MEAS_NAN;
    return (type == VIEKF_FEAT && E.id >= 0 && E.aux != 2) ? VIEKF_MEAS_NEW_FEATURE : VIEKF_MEAS_INVALID;
  }
  if (!E.active) return live ? VIEKF_MEAS_SUCCESS : VIEKF_MEAS_NAN;  // logged only (vi_ekf_meas.cpp:230)
  HReg H;
  if (!fast_model(s + O_X, p, type, rd, zv, -1, H)) {
    double* Hc = s + O_M;
    double* res = s + O_M + 18;
    int* cols = (int*)(s + O_M + 24);
    if (sub == 0) oct_generic_model(s, p, type, zdim, zv);
    __syncwarp();
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
  }
  const int cs = (type == VIEKF_ACC) ? (p.use_drag ? CS_ACC_DRAG : CS_ACC_GRAV)
               : (type == VIEKF_ALT) ? CS_ALT : (type == VIEKF_ATT) ? CS_ATT : (type == VIEKF_POS) ? CS_POS : (type == VIEKF_VEL) ? CS_VEL : CS_NONE;
  if (rd == 1) return oct_update_core<1>(s, p, E, sub, lane, F, Pa, Pb, H, cs, live);
  if (rd == 2) return oct_update_core<2>(s, p, E, sub, lane, F, Pa, Pb, H, cs, live);
  return oct_update_core<3>(s, p, E, sub, lane, F, Pa, Pb, H, cs, live);
}

// ------------------------------------------------------------------------------------------------------------------
// KFRESET / KEEP / STAT: rare events, evaluated on a packed copy of P in the exchange area
// ------------------------------------------------------------------------------------------------------------------
// keyframe_reset (vi_ekf_kfr.cpp:56-157) for the filters of the warp with `go` set
static __device__ __noinline__ void oct_kfreset_pose(double* s, const KParams& p, int64_t b) {
  Ctx<32> c;
  c.x = s + O_X;
  c.P = s + O_XA;
  c.b = b;
  kfreset_pose(c, p, s + O_M);
}
static __device__ __noinline__ void oct_kfreset_cov(double* s, int sub, bool go) {
  double* Pk = s + O_XA;
  const double* Nb = s + O_M;
  if (go) {
    // P = N P N^T with N = I except N(POS,POS) = 0 and N(ATT,ATT) = Nb, on the packed lower triangle
    for (int i = sub; i < 16; i += 8) {
      if (i < 3) {
        for (int j = 0; j <= i; j++) Pk[plo(i, j)] = 0.0;
      } else if (i < 6) {
        for (int j = 0; j < 3; j++) Pk[plo(i, j)] = 0.0;
        double v[3], o[3];
        for (int a = 0; a < 3; a++) v[a] = Pk[plo(6 + a, i)];
        for (int a = 0; a < 3; a++) o[a] = Nb[a * 3] * v[0] + Nb[a * 3 + 1] * v[1] + Nb[a * 3 + 2] * v[2];
        for (int a = 0; a < 3; a++) Pk[plo(6 + a, i)] = o[a];
      } else if (i == 6) {
        for (int r = 6; r < 9; r++)
          for (int j = 0; j < 3; j++) Pk[plo(r, j)] = 0.0;
        double A[9], Bm[9];
        for (int a = 0; a < 3; a++)
          for (int bb = 0; bb < 3; bb++) A[a * 3 + bb] = Pk[psym(6 + a, 6 + bb)];
        for (int a = 0; a < 3; a++)
          for (int bb = 0; bb < 3; bb++) Bm[a * 3 + bb] = Nb[a * 3] * A[bb] + Nb[a * 3 + 1] * A[3 + bb] + Nb[a * 3 + 2] * A[6 + bb];
        for (int a = 0; a < 3; a++)
          for (int bb = 0; bb <= a; bb++)
            Pk[plo(6 + a, 6 + bb)] = Bm[a * 3] * Nb[bb * 3] + Bm[a * 3 + 1] * Nb[bb * 3 + 1] + Bm[a * 3 + 2] * Nb[bb * 3 + 2];
      } else if (i >= 9) {
        for (int j = 0; j < 3; j++) Pk[plo(i, j)] = 0.0;
        double v[3];
        for (int a = 0; a < 3; a++) v[a] = Pk[plo(i, 6 + a)];
        for (int a = 0; a < 3; a++) Pk[plo(i, 6 + a)] = Nb[a * 3] * v[0] + Nb[a * 3 + 1] * v[1] + Nb[a * 3 + 2] * v[2];
      }
    }
  }
}
__device__ __forceinline__ void oct_kfreset(double* s, const KParams& p, int sub, int64_t b, bool go, double (&Pa)[16], double (&Pb)[16]) {
  double* Pk = s + O_XA;
  rows_to_packed(Pk, sub, Pa, Pb);
  __syncwarp();
  if (go && sub == 0) oct_kfreset_pose(s, p, b);
  __syncwarp();
  oct_kfreset_cov(s, sub, go);
  __syncwarp();
  if (go) rows_from_packed(Pk, sub, Pa, Pb);
  __syncwarp();
}

// keep_only_features (vi_ekf_feat.cpp:81-142) with nothing tracked: only the keyframe bookkeeping is left
__device__ __forceinline__ void oct_keep(double* s, const KParams& p, const viekf_event& E, int sub, int64_t b, bool valid, Flt& F,
                                         double (&Pa)[16], double (&Pb)[16]) {
  const int nf = E.n_ids;
  const int32_t* feats = p.id_pool + E.ids_off;
  const int nkf = F.nkf;
  const bool reset = valid && p.use_kfreset && nkf > 0 && (0.0 / (double)nkf < p.kf_thresh);
  if (__any_sync(0xffffffffu, reset)) oct_kfreset(s, p, sub, b, reset, Pa, Pb);
  if (p.use_kfreset && (reset || nkf == 0)) {
    const int m = nf < KF_CAP ? nf : KF_CAP;
    if (valid) {
      int32_t* kf = p.kf_ids + b * KF_CAP;
      for (int i = sub; i < m; i += 8) kf[i] = feats[i];
    }
    F.nkf = m;
  }
}

__device__ __forceinline__ bool oct_nan_scan(const double* s, int sub, const double (&Pa)[16], const double (&Pb)[16]) {
  bool bad = false;
#pragma unroll
  for (int j = 0; j < 16; j++) bad |= (Pa[j] != Pa[j]) || (Pb[j] != Pb[j]);
  const double* x = s + O_X;
  bad |= (x[sub] != x[sub]) || (x[sub + 8] != x[sub + 8]) || (x[16] != x[16]);
  return bad;
}

static __device__ __noinline__ void oct_stat_err(double* s, const KParams& p, const viekf_event& E, int64_t b, int64_t ginst) {
  {
  double* er = s + O_M;
    double xt[17];
    for (int i = 0; i < 8; i++) xt[i] = E.z[i];
    for (int i = 0; i < 9; i++) xt[8 + i] = E.R[i];
    if (E.noise >= 0 && p.values) {
      const double* v = p.values + b * p.vstride + 8 * (int64_t)E.noise;
      for (int i = 0; i < 6; i++) xt[10 + i] = v[i];
    }
    if (p.seed) {
      const double* bi = s + O_BIAS;
      Q4 qbu = ldq(p.qbu);
      V3 ba = rota(qbu, mk3(bi[0], bi[1], bi[2])), bg = rota(qbu, mk3(bi[3], bi[4], bi[5]));
      xt[10] += ba.x; xt[11] += ba.y; xt[12] += ba.z;
      xt[13] += bg.x; xt[14] += bg.y; xt[15] += bg.z;
    }
    {  // the global truth in the frame of the filter's current node (inverse of get_global_pose, vi_ekf_kfr.cpp:14-22)
      const FilterMeta& M = p.meta[b];
      const Q4 nq = ldq(M.node_q);
      const V3 pr = rotp(nq, mk3(xt[0] - M.node_t[0], xt[1] - M.node_t[1], xt[2] - M.node_t[2]));
      const Q4 qr = otimes(Q4{nq.w, -nq.x, -nq.y, -nq.z}, ldq(xt + 6));
      xt[0] = pr.x; xt[1] = pr.y; xt[2] = pr.z;
      stq(xt + 6, qr);
    }
    const double* x = s + O_X;
    for (int i = 0; i < 6; i++) er[i] = xt[i] - x[i];
    V3 d = qboxminus(ldq(xt + 6), ldq(x + 6));
    er[6] = d.x; er[7] = d.y; er[8] = d.z;
    for (int i = 0; i < 7; i++) er[9 + i] = xt[10 + i] - x[10 + i];
  }
}
// e^T B^-1 e for the 16 x 16 base block by the eight lanes of the filter: nees_warp<16> of viekf_dev.cuh (right-looking Cholesky,
// same operation order per element) with the rows `sub` and `sub + 8` on each lane and 8-lane shuffles.
static __device__ __forceinline__ double oct_nees16(const double* Pk, const double* er, int sub) {
  double aa[8], ab[16];
#pragma unroll
  for (int j = 0; j < 8; j++) aa[j] = (j <= sub) ? Pk[plo(sub, j <= sub ? j : 0)] : 0.0;
#pragma unroll
  for (int j = 0; j < 16; j++) ab[j] = (j <= sub + 8) ? Pk[plo(sub + 8, j <= sub + 8 ? j : 0)] : 0.0;
  double ra = er[sub], rb = er[sub + 8];
  double q = 0;
  bool bad = false;
#pragma unroll
  for (int j = 0; j < 16; j++) {
    const double sjj = __shfl_sync(0xffffffffu, (j < 8) ? aa[j < 8 ? j : 0] : ab[j], j & 7, 8);
    bad |= !(sjj > 0.0);
    const double ljj = sqrt(sjj);
    double la = 0.0, lb;
    if (j < 8) la = (sub == j) ? ljj : aa[j < 8 ? j : 0] / ljj;
    lb = (sub + 8 == j) ? ljj : ab[j] / ljj;
    const double yj = __shfl_sync(0xffffffffu, ((j < 8) ? ra : rb) / ljj, j & 7, 8);
    q += yj * yj;
    if (j < 8 && sub > j) ra -= la * yj;
    if (sub + 8 > j) rb -= lb * yj;
#pragma unroll
    for (int k = j + 1; k < 16; k++) {
      const double lkj = __shfl_sync(0xffffffffu, (k < 8) ? la : lb, k & 7, 8);
      if (k < 8 && sub >= k) aa[k < 8 ? k : 0] -= la * lkj;
      if (sub + 8 >= k) ab[k] -= lb * lkj;
    }
  }
  return bad ? nan("") : q;
}
static __device__ __noinline__ void oct_stat_nees(double* s, int sub) {
  const double q16 = oct_nees16(s + O_XA, s + O_M, sub);
  const double q6 = nees_warp<6, 8>(s + O_XA, s + O_M, sub);
  if (sub == 0) {
    s[O_M + 16] = q16;
    s[O_M + 17] = q6;
  }
}
// STAT (same sums as ev_stat of viekf_dev.cuh; the NEES factorisations run one filter at a time on the whole warp)
__device__ __forceinline__ void oct_stat(double* s, double* sw, const KParams& p, const viekf_event& E, int sub, int lane, int64_t b,
                                         int64_t ginst, bool valid, Flt& F, const double (&Pa)[16], const double (&Pb)[16]) {
  if (!p.stats || E.id < 0 || E.id >= p.n_buckets) return;
  double* st = p.stats + (size_t)E.id * VIEKF_STAT_WIDTH;
  const unsigned badm = __ballot_sync(0xffffffffu, oct_nan_scan(s, sub, Pa, Pb)) & octet_mask(lane);
  if (badm) F.flags |= VIEKF_FLAG_NAN_STATE;
  const bool healthy = !(F.flags & VIEKF_FLAG_NAN_STATE);
  if (!healthy && valid && sub == 0) atomicAdd(st + 38, 1.0);
  double* Pk = s + O_XA;
  double* er = s + O_M;
  double* out = s + O_M + 16;
  rows_to_packed(Pk, sub, Pa, Pb);
  if (sub == 0) oct_stat_err(s, p, E, b, ginst);
  __syncwarp();
  oct_stat_nees(s, sub);
  __syncwarp();
  if (healthy && valid) {
    const bool ok16 = isfinite(out[0]), ok6 = isfinite(out[1]);
    for (int k = sub; k < VIEKF_STAT_WIDTH; k += 8) {
      double v;
      if (k == 0) v = 1.0;
      else if (k <= 16) v = er[k - 1];
      else if (k <= 32) v = er[k - 17] * er[k - 17];
      else if (k == 33) v = ok16 ? out[0] : 0.0;
      else if (k == 34) v = ok6 ? out[1] : 0.0;
      else if (k == 35) v = (double)F.ngated;
      else if (k == 36) v = (double)p.meta[b].n_fix;
      else if (k == 37) v = ok16 ? 1.0 : 0.0;
      else if (k == 39) v = ok6 ? 1.0 : 0.0;
      else continue;
      if (v != 0.0 || k <= 32) atomicAdd(st + k, v);
    }
  }
  __syncwarp();
}

// ------------------------------------------------------------------------------------------------------------------
// the persistent tape executor
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32 * VIEKF_OCTET_WPC, VIEKF_OCTET_MINB) viekf_octet_kernel(const __grid_constant__ KParams p) {
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane & 7, fo = lane >> 3;
  double* sw = smem + (size_t)(warp * 4) * O_TOTAL;
  double* s = sw + (size_t)fo * O_TOTAL;
  double* evb = smem + (size_t)(VIEKF_OCTET_WPC * 4) * O_TOTAL + warp * 64;  // two staged viekf_event records (31 doubles each) per warp
  static_assert(sizeof(viekf_event) == 31 * sizeof(double), "event staging assumes 31 doubles per record");
  // the trip count is uniform over the CTA (lock-step barrier); a warp or lane without a filter runs on a copy of the last one
  for (int64_t q0 = (int64_t)blockIdx.x * (VIEKF_OCTET_WPC * 4); q0 < p.B; q0 += (int64_t)gridDim.x * (VIEKF_OCTET_WPC * 4)) {
    int64_t b = q0 + warp * 4 + fo;
    const bool valid = b < p.B;  // lanes without a filter still take part in every __syncwarp
    if (!valid) b = p.B - 1;
    const int64_t ginst = p.inst_off + b;
    FilterMeta& M = p.meta[b];
    // ---- load the filter ----
    Flt F;
    F.t = M.t;
    F.started = M.started;
    F.nkf = M.n_kf;
    F.flags = (int)M.flags;
    F.ngated = (int)M.n_gated;
    F.nnan = (int)M.n_nan_gain;
    double Pa[16], Pb[16];
    __syncwarp();
    {
      const double* gx = p.x + b * p.xs;
      s[O_X + sub] = gx[sub];
      s[O_X + sub + 8] = gx[sub + 8];
      if (sub == 0) s[O_X + 16] = gx[16];
      if (sub < 6) {
        s[O_UB + sub] = M.ub[sub];
        s[O_BIAS + sub] = (p.seed && p.bias_range[sub] != 0.0) ? p.bias_range[sub] * viekf_uniform_pm1(p.seed, (uint64_t)ginst, VIEKF_NOISE_EVENT_BIAS, sub) : 0.0;
      }
      rows_from_packed(p.P + b * p.ps, sub, Pa, Pb);
      if (p.ev_count > 0 && lane < 31) evb[lane] = ((const double*)(p.ev + p.ev_first))[lane];
    }
    __syncwarp();
    // ---- run the tape slice ----
    for (int64_t rel = 0; rel < p.ev_count; rel++) {
      const int64_t e = p.ev_first + rel;
      const viekf_event& E = *(const viekf_event*)(evb + 32 * (int)(rel & 1));
      // the next record travels through a register while this event runs and is parked in the other slot at the end
      const bool more = rel + 1 < p.ev_count && lane < 31;
      double nxt = 0.0;
      if (more) nxt = ((const double*)(p.ev + e + 1))[lane];
      const int kind = E.kind;
      if (kind == VIEKF_EV_PROP || kind == VIEKF_EV_UPD) {
        oct_resolve(p, E, e, b, ginst, sub, s);
        __syncwarp();
      }
      int result = VIEKF_MEAS_NONE;
      switch (kind) {
        case VIEKF_EV_PROP:
          oct_propagate(s, p, E, sub, F, Pa, Pb);
          break;
        case VIEKF_EV_UPD:
          result = oct_update(s, p, E, sub, lane, F, Pa, Pb);
          break;
        case VIEKF_EV_INIT:
          result = 0;  // init_feature refuses: no slot (vi_ekf_feat.cpp:9)
          break;
        case VIEKF_EV_KEEP:
          oct_keep(s, p, E, sub, b, valid, F, Pa, Pb);
          break;
        case VIEKF_EV_KFRESET:
          oct_kfreset(s, p, sub, b, valid, Pa, Pb);
          break;
        case VIEKF_EV_STAT:
          oct_stat(s, sw, p, E, sub, lane, b, ginst, valid, F, Pa, Pb);
          break;
        default:  // CLEAR: nothing is tracked
          break;
      }
      if (more) evb[32 * (int)((rel + 1) & 1) + lane] = nxt;
#if VIEKF_OCTET_LOCKSTEP
      __syncthreads();
#else
      __syncwarp();
#endif
      if (p.results && sub == 0 && valid && b >= p.res_first && b < p.res_first + p.res_count)
        p.results[(b - p.res_first) * p.n_events + e] = (uint8_t)result;
    }
    // ---- store the filter; NaNsInTheHouse at the end of the slice ----
    {
      const unsigned badm = __ballot_sync(0xffffffffu, oct_nan_scan(s, sub, Pa, Pb)) & octet_mask(lane);
      if (badm) F.flags |= VIEKF_FLAG_NAN_STATE;
    }
    if (valid) {
      if (sub == 0) {
        M.t = F.t;
        M.started = F.started;
        M.n_kf = F.nkf;
        M.flags = (uint32_t)F.flags;
        M.n_gated = (uint32_t)F.ngated;
        M.n_nan_gain = (uint32_t)F.nnan;
      }
      if (sub < 6) M.ub[sub] = s[O_UB + sub];
      double* gx = p.x + b * p.xs;
      gx[sub] = s[O_X + sub];
      gx[sub + 8] = s[O_X + sub + 8];
      if (sub == 0) gx[16] = s[O_X + 16];
      rows_to_packed(p.P + b * p.ps, sub, Pa, Pb);
    }
    __syncwarp();
  }
}

size_t octet_smem_bytes() { return ((size_t)O_TOTAL * 4 + 64) * VIEKF_OCTET_WPC * sizeof(double); }

cudaError_t launch_octet(const KParams& p, int sm_count, cudaStream_t stream) {
  if (p.N != 0) return cudaErrorInvalidValue;
  const size_t smem = octet_smem_bytes();
  cudaError_t e = cudaFuncSetAttribute(viekf_octet_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(viekf_octet_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return e;
  int occ = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, viekf_octet_kernel, 32 * VIEKF_OCTET_WPC, smem);
  if (e != cudaSuccess) return e;
  if (occ < 1) return cudaErrorInvalidConfiguration;
  const int64_t need = (p.B + 4 * VIEKF_OCTET_WPC - 1) / (4 * VIEKF_OCTET_WPC);
  int64_t grid = (int64_t)sm_count * occ;  // persistent: a whole number of waves of the SM count
  if (need < grid) grid = need;
  if (grid < 1) grid = 1;
  viekf_octet_kernel<<<(unsigned)grid, 32 * VIEKF_OCTET_WPC, smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace vk
