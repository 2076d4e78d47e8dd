// viekf_oracle.cpp -- TEST INFRASTRUCTURE (see header).  Dense restatement of reference src/vi_ekf/*.cpp.
#include "viekf_oracle.hpp"
#include "../include/viekf_noise.h"
#include <array>
#include <cstdio>

namespace viekf_oracle {

enum { xPOS = 0, xVEL = 3, xATT = 6, xB_A = 10, xB_G = 13, xMU = 16, xZ = 17 };
enum { dxPOS = 0, dxVEL = 3, dxATT = 6, dxB_A = 9, dxB_G = 12, dxMU = 15, dxZ = 16 };
enum { uA = 0, uG = 3 };
static const V3 gravity = {{0, 0, 9.80665}};  // include/vi_ekf.h:70-74

// ---- dense products (column-major, k ascending) -------------------------------------------------
static void gemm(const Mat& A, const Mat& B, Mat& C) {  // C = A*B
  C.zero();
  const int n = A.r, K = A.c, m = B.c;
  for (int j = 0; j < m; j++)
    for (int k = 0; k < K; k++) {
      const double b = B(k, j);
      const double* a = &A.d[(size_t)k * n];
      double* c = &C.d[(size_t)j * n];
      for (int i = 0; i < n; i++) c[i] += a[i] * b;
    }
}
static void gemm_nt(const Mat& A, const Mat& B, Mat& C) {  // C = A*B^T
  C.zero();
  const int n = A.r, K = A.c, m = B.r;
  for (int j = 0; j < m; j++)
    for (int k = 0; k < K; k++) {
      const double b = B(j, k);
      const double* a = &A.d[(size_t)k * n];
      double* c = &C.d[(size_t)j * n];
      for (int i = 0; i < n; i++) c[i] += a[i] * b;
    }
}
static void set_block(Mat& M, int r0, int c0, const M3& b) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) M(r0 + i, c0 + j) = b.m[i][j];
}

// Partial-pivot LU inverse of an r x r (r<=3) column-major matrix -- what Eigen's dynamic-size
// .inverse() does (vi_ekf_meas.cpp:232).
static void small_inverse(const double* S, int r, double* Si) {
  double a[3][6];
  for (int i = 0; i < r; i++)
    for (int j = 0; j < r; j++) {
      a[i][j] = S[i + j * r];
      a[i][j + r] = (i == j) ? 1.0 : 0.0;
    }
  for (int c = 0; c < r; c++) {
    int p = c;
    for (int i = c + 1; i < r; i++)
      if (std::fabs(a[i][c]) > std::fabs(a[p][c])) p = i;
    if (p != c)
      for (int j = 0; j < 2 * r; j++) std::swap(a[c][j], a[p][j]);
    for (int i = c + 1; i < r; i++) {
      double f = a[i][c] / a[c][c];
      for (int j = c; j < 2 * r; j++) a[i][j] -= f * a[c][j];
    }
  }
  for (int c = r - 1; c >= 0; c--) {
    for (int j = 0; j < r; j++) {
      double s = a[c][r + j];
      for (int k = c + 1; k < r; k++) s -= a[c][k] * a[k][r + j];
      a[c][r + j] = s / a[c][c];
    }
  }
  for (int i = 0; i < r; i++)
    for (int j = 0; j < r; j++) Si[i + j * r] = a[i][r + j];
}

// ---- construction: VIEKF::init() + load()/init(...) (vi_ekf.cpp:18-155) ---------------------------
VIEKF::VIEKF(const viekf_config& cfg, int NF_, int hist)
    : NF(NF_), MAX_X(17 + 5 * NF_), MAX_DX(16 + 3 * NF_), HIST(hist), cfg_(cfg) {
  x_.assign(HIST, std::vector<double>(MAX_X, 0.0));
  P_.assign(HIST, Mat(MAX_DX, MAX_DX));
  t_.assign(HIST, NAN);
  xp_.assign(MAX_X, 0.0);
  dx_.assign(MAX_DX, 0.0);
  lambda_.assign(MAX_DX, 0.0);
  Qx_ = Mat(MAX_DX, MAX_DX);
  Lambda_ = Mat(MAX_DX, MAX_DX);
  A_ = Mat(MAX_DX, MAX_DX);
  G_ = Mat(MAX_DX, 6);
  T1_ = Mat(MAX_DX, MAX_DX);
  T2_ = Mat(MAX_DX, MAX_DX);
  T3_ = Mat(MAX_DX, MAX_DX);
  H_ = Mat(3, MAX_DX);
  K_ = Mat(MAX_DX, 3);
  for (int i = 0; i < 4; i++) zhat_[i] = 0;
  for (int i = 0; i < 36; i++) global_cov_[i] = 0;

  for (int i = 0; i < 17; i++) x_[i_][i] = cfg.x0[i];
  for (int i = 0; i < 16; i++) {
    P_[i_](i, i) = cfg.P0[i];
    Qx_(i, i) = cfg.Qx[i];
    lambda_[i] = cfg.lambda[i];
  }
  for (int i = 0; i < 6; i++) Qu_[i] = cfg.Qu[i];
  for (int i = 0; i < 3; i++) P0_feat_[i] = cfg.P0_feat[i];  // set on both init paths
  for (int f = 0; f < NF; f++)
    for (int k = 0; k < 3; k++) {
      P_[i_](dxZ + 3 * f + k, dxZ + 3 * f + k) = cfg.P0_feat[k];
      Qx_(dxZ + 3 * f + k, dxZ + 3 * f + k) = cfg.Qx_feat[k];
      lambda_[dxZ + 3 * f + k] = cfg.lambda_feat[k];
    }
  // Lambda_ = 1 lambda^T + lambda 1^T - lambda lambda^T   (vi_ekf.cpp:83,146)
  for (int j = 0; j < MAX_DX; j++)
    for (int i = 0; i < MAX_DX; i++) Lambda_(i, j) = (1.0 * lambda_[j] + lambda_[i] * 1.0) - lambda_[i] * lambda_[j];
  cam_F_[0] = cfg.focal_len[0];
  cam_F_[1] = cfg.focal_len[1];
  q_b_c_ = Quat(cfg.q_b_c);
  q_b_u_ = Quat(cfg.q_b_u);
  p_b_c_ = v3(cfg.p_b_c[0], cfg.p_b_c[1], cfg.p_b_c[2]);
}

// ---- vi_ekf_helper.cpp:88-111 --------------------------------------------------------------------
void VIEKF::boxplus(const std::vector<double>& x, const std::vector<double>& dx, std::vector<double>& out) const {
  for (int i = 0; i < 6; i++) out[xPOS + i] = x[xPOS + i] + dx[dxPOS + i];
  Quat q = Quat(&x[xATT]).boxplus(v3(dx[dxATT], dx[dxATT + 1], dx[dxATT + 2]));
  for (int i = 0; i < 4; i++) out[xATT + i] = q.q[i];
  for (int i = 0; i < 7; i++) out[xB_A + i] = x[xB_A + i] + dx[dxB_A + i];
  for (int i = 0; i < len_features_; i++) {
    Quat qz = q_feat_boxplus(Quat(&x[xZ + i * 5]), &dx[dxZ + 3 * i]);
    for (int k = 0; k < 4; k++) out[xZ + i * 5 + k] = qz.q[k];
    out[xZ + i * 5 + 4] = x[xZ + i * 5 + 4] + dx[dxZ + 3 * i + 2];
  }
}

void VIEKF::boxminus(const std::vector<double>& x1, const std::vector<double>& x2, std::vector<double>& out) const {
  for (int i = 0; i < 6; i++) out[dxPOS + i] = x1[xPOS + i] - x2[xPOS + i];
  V3 d = Quat(&x1[xATT]).boxminus(Quat(&x2[xATT]));
  for (int i = 0; i < 3; i++) out[dxATT + i] = d[i];
  for (int i = 0; i < 7; i++) out[dxB_A + i] = x1[xB_A + i] - x2[xB_A + i];
  for (int i = 0; i < len_features_; i++) {
    q_feat_boxminus(Quat(&x1[xZ + i * 5]), Quat(&x2[xZ + i * 5]), &out[dxZ + i * 3]);
    out[dxZ + i * 3 + 2] = x1[xZ + i * 5 + 4] - x2[xZ + i * 5 + 4];
  }
}

int VIEKF::global_to_local_feature_id(int id) const {
  for (size_t i = 0; i < current_feature_ids_.size(); i++)
    if (current_feature_ids_[i] == id) return (int)i;
  return -1;
}

// ---- vi_ekf_error.cpp:6-19 -------------------------------------------------------------------------
// NaNsInTheHouse(): any NaN in the live part of x or in the live block of P.  The reference only evaluates it inside the
// debug-build NAN_CHECK macro (include/vi_ekf.h:29-33); here it latches the sticky VIEKF_FLAG_NAN_STATE.
bool VIEKF::nans_in_the_house() const {
  const int x_max = xZ + len_features_ * 5, dx_max = dxZ + len_features_ * 3;
  const std::vector<double>& x = x_[i_];
  const Mat& P = P_[i_];
  for (int i = 0; i < x_max; i++)
    if (x[i] != x[i]) return true;
  for (int j = 0; j < dx_max; j++)
    for (int i = 0; i < dx_max; i++)
      if (P(i, j) != P(i, j)) return true;
  return false;
}
bool VIEKF::check_nan_state() {
  if (nans_in_the_house()) flags |= VIEKF_FLAG_NAN_STATE;
  return (flags & VIEKF_FLAG_NAN_STATE) != 0;
}

// ---- vi_ekf_helper.cpp:128-156 ---------------------------------------------------------------------
void VIEKF::fix_depth() {
  std::vector<double>& x = x_[i_];
  Mat& P = P_[i_];
  const double reset = 1.0 / (2.0 * cfg_.min_depth);
  for (int i = 0; i < len_features_; i++) {
    int xR = xZ + 5 * i + 4, dxR = dxZ + 3 * i + 2;
    if (x[xR] != x[xR]) {
      x[xR] = reset;
      flags |= VIEKF_FLAG_DEPTH_NAN_FIXED;
      n_fix++;
    }
    min_depth_margin = std::min(min_depth_margin, std::min(std::fabs(x[xR]), std::fabs(x[xR] - 1e2)));
    if (x[xR] < 0.0) {
      double err = reset - x[xR];
      P(dxR, dxR) += err * err;
      x[xR] = reset;
      flags |= VIEKF_FLAG_DEPTH_NEG_FIXED;
      n_fix++;
    } else if (x[xR] > 1e2) {
      P(dxR, dxR) = P0_feat_[2];
      x[xR] = reset;
      flags |= VIEKF_FLAG_DEPTH_BIG_FIXED;
      n_fix++;
    }
  }
}

// ---- vi_ekf_dyn.cpp:14-135 --------------------------------------------------------------------------
void VIEKF::dynamics(const std::vector<double>& x, const double u[6], bool state, bool jac) {
  if (state) std::fill(dx_.begin(), dx_.end(), 0.0);
  if (jac) {
    A_.zero();
    G_.zero();
  }
  V3 vel = v3(x[xVEL], x[xVEL + 1], x[xVEL + 2]);
  Quat q_I_b(&x[xATT]);
  V3 acc = v3(u[uA] - x[xB_A], u[uA + 1] - x[xB_A + 1], u[uA + 2] - x[xB_A + 2]);
  V3 omega = v3(u[uG] - x[xB_G], u[uG + 1] - x[xB_G + 1], u[uG + 2] - x[xB_G + 2]);
  V3 acc_z = v3(0, 0, acc[2]);
  double mu = x[xMU];
  M3 R_I_b = q_I_b.R();
  V3 gravity_B = q_I_b.rotp(gravity);
  V3 vel_xy = v3(vel[0], vel[1], 0.0);
  const bool drag = cfg_.use_drag_term != 0;

  if (state) {
    V3 pd = q_I_b.rota(vel);
    V3 vd;
    if (drag)
      vd = acc_z + gravity_B - cross(omega, vel) - mu * vel_xy;
    else
      vd = acc + gravity_B - cross(omega, vel);
    for (int i = 0; i < 3; i++) {
      dx_[dxPOS + i] = pd[i];
      dx_[dxVEL + i] = vd[i];
      dx_[dxATT + i] = omega[i];
    }
  }
  if (jac) {
    M3 Rt = transpose(R_I_b);
    set_block(A_, dxPOS, dxVEL, Rt);
    set_block(A_, dxPOS, dxATT, mul(scale(Rt, -1.0), skew(vel)));
    if (drag) {
      M3 D = {{{1, 0, 0}, {0, 1, 0}, {0, 0, 0}}};  // I_2x3^T I_2x3
      set_block(A_, dxVEL, dxVEL, sub(scale(D, -mu), skew(omega)));
      M3 Z = {{{0, 0, 0}, {0, 0, 0}, {0, 0, -1}}};
      set_block(A_, dxVEL, dxB_A, Z);
      for (int i = 0; i < 3; i++) A_(dxVEL + i, dxMU) = -vel_xy[i];
    } else {
      set_block(A_, dxVEL, dxVEL, scale(skew(omega), -1.0));
      set_block(A_, dxVEL, dxB_A, scale(eye3(), -1.0));
    }
    set_block(A_, dxVEL, dxATT, skew(gravity_B));
    set_block(A_, dxVEL, dxB_G, scale(skew(vel), -1.0));
    set_block(A_, dxATT, dxATT, scale(skew(omega), -1.0));
    set_block(A_, dxATT, dxB_G, scale(eye3(), -1.0));
    if (drag) {
      M3 Z = {{{0, 0, 0}, {0, 0, 0}, {0, 0, -1}}};
      set_block(G_, dxVEL, uA, Z);
    } else {
      set_block(G_, dxVEL, uA, scale(eye3(), -1.0));
    }
    set_block(G_, dxVEL, uG, scale(skew(vel), -1.0));
    set_block(G_, dxATT, uG, scale(eye3(), -1.0));
  }

  // camera dynamics (vi_ekf_dyn.cpp:83-95)
  V3 vel_c_i = q_b_c_.rotp(vel + cross(omega, p_b_c_));
  V3 omega_c_i = q_b_c_.rotp(omega);
  M3 skew_vel_c = skew(vel_c_i);
  M3 skew_p_b_c = skew(p_b_c_);
  M3 R_b_c = q_b_c_.R();

  for (int i = 0; i < len_features_; i++) {
    int xZETA_i = xZ + i * 5, xRHO_i = xZ + 5 * i + 4, dxZETA_i = dxZ + i * 3, dxRHO_i = dxZ + i * 3 + 2;
    Quat q_zeta(&x[xZETA_i]);
    double rho = x[xRHO_i];
    V3 zeta = q_zeta.rota(v3(0, 0, 1));
    M32 T_z = T_zeta(q_zeta);
    M3 skew_zeta = skew(zeta);
    double rho2 = rho * rho;
    // helper: T_z^T * (3x3) -> 2x3, T_z^T * v -> 2
    auto Tt_v = [&](const V3& v, double o[2]) {
      for (int a = 0; a < 2; a++) o[a] = T_z.m[0][a] * v[0] + T_z.m[1][a] * v[1] + T_z.m[2][a] * v[2];
    };
    auto Tt_M = [&](const M3& M, double o[2][3]) {
      for (int a = 0; a < 2; a++)
        for (int c = 0; c < 3; c++) o[a][c] = T_z.m[0][a] * M.m[0][c] + T_z.m[1][a] * M.m[1][c] + T_z.m[2][a] * M.m[2][c];
    };
    if (state) {
      double o[2];
      Tt_v(omega_c_i + rho * cross(zeta, vel_c_i), o);
      dx_[dxZETA_i] = -o[0];
      dx_[dxZETA_i + 1] = -o[1];
      dx_[dxRHO_i] = rho2 * dot(zeta, vel_c_i);
    }
    if (jac) {
      double o[2][3];
      // (zeta, V) = -rho T^T [zeta]x R_bc
      Tt_M(mul(skew_zeta, R_b_c), o);
      for (int a = 0; a < 2; a++)
        for (int c = 0; c < 3; c++) A_(dxZETA_i + a, dxVEL + c) = -rho * o[a][c];
      // (zeta, b_g) = -T^T (rho [zeta]x R_bc [p_bc]x - R_bc)
      M3 Mbg = sub(scale(mul(mul(skew_zeta, R_b_c), skew_p_b_c), rho), R_b_c);
      Tt_M(Mbg, o);
      for (int a = 0; a < 2; a++)
        for (int c = 0; c < 3; c++) {
          A_(dxZETA_i + a, dxB_G + c) = -o[a][c];
          G_(dxZETA_i + a, uG + c) = -o[a][c];
        }
      // (zeta, zeta) = -T^T ([omega_c + rho zeta x v_c]x + rho [v_c]x [zeta]x) T
      M3 Mzz = add(skew(omega_c_i + rho * cross(zeta, vel_c_i)), scale(mul(skew_vel_c, skew_zeta), rho));
      Tt_M(Mzz, o);
      for (int a = 0; a < 2; a++)
        for (int b = 0; b < 2; b++)
          A_(dxZETA_i + a, dxZETA_i + b) = -(o[a][0] * T_z.m[0][b] + o[a][1] * T_z.m[1][b] + o[a][2] * T_z.m[2][b]);
      // (zeta, rho) = -T^T (zeta x v_c)
      double o2[2];
      Tt_v(cross(zeta, vel_c_i), o2);
      A_(dxZETA_i, dxRHO_i) = -o2[0];
      A_(dxZETA_i + 1, dxRHO_i) = -o2[1];
      // (rho, V) = rho^2 zeta^T R_bc ; (rho, b_g) = rho^2 zeta^T R_bc [p_bc]x
      M3 RP = mul(R_b_c, skew_p_b_c);
      for (int c = 0; c < 3; c++) {
        double zr = zeta[0] * R_b_c.m[0][c] + zeta[1] * R_b_c.m[1][c] + zeta[2] * R_b_c.m[2][c];
        double zrp = zeta[0] * RP.m[0][c] + zeta[1] * RP.m[1][c] + zeta[2] * RP.m[2][c];
        A_(dxRHO_i, dxVEL + c) = rho2 * zr;
        A_(dxRHO_i, dxB_G + c) = rho2 * zrp;
        G_(dxRHO_i, uG + c) = rho2 * zrp;
      }
      // (rho, zeta) = rho^2 zeta^T [v_c]x T
      double zs[3];
      for (int c = 0; c < 3; c++) zs[c] = zeta[0] * skew_vel_c.m[0][c] + zeta[1] * skew_vel_c.m[1][c] + zeta[2] * skew_vel_c.m[2][c];
      for (int b = 0; b < 2; b++) A_(dxRHO_i, dxZETA_i + b) = rho2 * (zs[0] * T_z.m[0][b] + zs[1] * T_z.m[1][b] + zs[2] * T_z.m[2][b]);
      // (rho, rho) = 2 rho zeta . v_c
      A_(dxRHO_i, dxRHO_i) = 2 * rho * dot(zeta, vel_c_i);
    }
  }
}

// ---- vi_ekf.cpp:262-318 --------------------------------------------------------------------------------
void VIEKF::propagate_state(const double u[6], double t, bool save_input) {
  double ub[6];
  V3 a = q_b_u_.rota(v3(u[0], u[1], u[2])), g = q_b_u_.rota(v3(u[3], u[4], u[5]));
  for (int i = 0; i < 3; i++) {
    ub[i] = a[i];
    ub[3 + i] = g[i];
  }
  for (int i = 0; i < 6; i++) ub_last_[i] = ub[i];
  if (save_input) {
    std::array<double, 6> s;
    for (int i = 0; i < 6; i++) s[i] = ub[i];
    u_.push_front({t, s});
  }
  if (std::isnan(start_t_)) {
    start_t_ = t;
    t_[i_] = t;
    return;
  }
  double dt = t - t_[i_];
  if (std::fabs(dt) < 1e-6) return;
  if (dt < 0) {
    flags |= VIEKF_FLAG_BACKWARD_DT;
    return;
  }
  dynamics(x_[i_], ub, true, true);
  int ip = (i_ + 1) % HIST;
  const int n = MAX_DX;

  // boxplus(x, dx*dt, x_[ip])  (vi_ekf.cpp:301)
  std::vector<double> dxdt(n);
  for (int i = 0; i < n; i++) dxdt[i] = dx_[i] * dt;
  boxplus(x_[i_], dxdt, x_[ip]);

  // G_ = (I + A dt/2 + A A dt dt/6) G dt ; A_ = I + A dt + A A dt dt/2   (vi_ekf.cpp:302-303)
  Mat& AA = T1_;
  gemm(A_, A_, AA);
  Mat& M = T2_;
  for (int j = 0; j < n; j++)
    for (int i = 0; i < n; i++) M(i, j) = ((i == j ? 1.0 : 0.0) + A_(i, j) * dt / 2.0) + AA(i, j) * dt * dt / 6.0;
  Mat Gd(n, 6);
  gemm(M, G_, Gd);
  for (auto& v : Gd.d) v *= dt;
  gemm(A_, A_, AA);  // evaluated twice in the reference; same values
  for (int j = 0; j < n; j++)
    for (int i = 0; i < n; i++) T2_(i, j) = ((i == j ? 1.0 : 0.0) + A_(i, j) * dt) + AA(i, j) * dt * dt / 2.0;
  std::swap(A_.d, T2_.d);  // A_ = discrete A
  G_ = Gd;

  // P_[ip] = A P A^T + G Qu G^T + Qx   (vi_ekf.cpp:304)
  gemm(A_, P_[i_], T1_);
  gemm_nt(T1_, A_, T3_);
  Mat GQ(n, 6);
  for (int k = 0; k < 6; k++)
    for (int i = 0; i < n; i++) GQ(i, k) = G_(i, k) * Qu_[k];
  Mat GQG(n, n);
  gemm_nt(GQ, G_, GQG);
  Mat& Pn = P_[ip];
  for (size_t k = 0; k < Pn.d.size(); k++) Pn.d[k] = (T3_.d[k] + GQG.d[k]) + Qx_.d[k];
  t_[ip] = t;
  i_ = ip;
  fix_depth();
}

// ---- measurement models: vi_ekf_meas.cpp:281-395 --------------------------------------------------------
void VIEKF::h(int type, const std::vector<double>& x, double zh[4], Mat& H, int id) const {
  H.zero();
  switch (type) {
    case VIEKF_ACC: {
      if (cfg_.use_drag_term) {
        double mu = x[xMU];
        for (int a = 0; a < 2; a++) {
          zh[a] = -mu * x[xVEL + a] + x[xB_A + a];
          H(a, dxVEL + a) = -mu;
          H(a, dxB_A + a) = 1.0;
          H(a, dxMU) = -x[xVEL + a];
        }
      } else {
        V3 gB = Quat(&x[xATT]).rotp(gravity);
        for (int a = 0; a < 3; a++) zh[a] = x[xB_A + a] - gB[a];
        M3 s = skew(-1.0 * gB);
        for (int a = 0; a < 3; a++)
          for (int c = 0; c < 3; c++) H(a, dxATT + c) = s.m[a][c];
        for (int a = 0; a < 3; a++) H(a, dxB_A + a) = 1.0;
      }
      break;
    }
    case VIEKF_ALT:
      zh[0] = -x[xPOS + 2];
      H(0, dxPOS + 2) = -1.0;
      break;
    case VIEKF_ATT:
      for (int a = 0; a < 4; a++) zh[a] = x[xATT + a];
      for (int a = 0; a < 3; a++) H(a, dxATT + a) = 1.0;
      break;
    case VIEKF_POS:
      for (int a = 0; a < 3; a++) {
        zh[a] = x[xPOS + a];
        H(a, xPOS + a) = 1.0;
      }
      break;
    case VIEKF_VEL:
      for (int a = 0; a < 3; a++) {
        zh[a] = x[xVEL + a];
        H(a, dxVEL + a) = 1.0;
      }
      break;
    case VIEKF_QZETA: {
      int i = global_to_local_feature_id(id);
      for (int a = 0; a < 4; a++) zh[a] = x[xZ + i * 5 + a];
      H(0, dxZ + i * 3) = 1.0;
      H(1, dxZ + i * 3 + 1) = 1.0;
      break;
    }
    case VIEKF_FEAT: {
      int i = global_to_local_feature_id(id);
      Quat q_zeta(&x[xZ + i * 5]);
      V3 zeta = q_zeta.rota(v3(0, 0, 1));
      M3 sk = skew(zeta);
      double ez = zeta[2];
      M32 T = T_zeta(q_zeta);
      zh[0] = cam_F_[0] * zeta[0] / ez + cfg_.cam_center[0];
      zh[1] = cam_F_[1] * zeta[1] / ez + cfg_.cam_center[1];
      // H = (1/ez) cam_F ((zeta e_z^T)/ez - I) [zeta]x T_z     (vi_ekf_meas.cpp:366)
      M3 B;
      for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) B.m[r][c] = ((c == 2) ? zeta[r] * 1.0 : 0.0) / ez - (r == c ? 1.0 : 0.0);
      double FB[2][3];
      for (int c = 0; c < 3; c++) {
        FB[0][c] = (1.0 / ez) * cam_F_[0] * B.m[0][c];
        FB[1][c] = (1.0 / ez) * cam_F_[1] * B.m[1][c];
      }
      double FBS[2][3];
      for (int a = 0; a < 2; a++)
        for (int c = 0; c < 3; c++) FBS[a][c] = FB[a][0] * sk.m[0][c] + FB[a][1] * sk.m[1][c] + FB[a][2] * sk.m[2][c];
      for (int a = 0; a < 2; a++)
        for (int b = 0; b < 2; b++) H(a, dxZ + i * 3 + b) = FBS[a][0] * T.m[0][b] + FBS[a][1] * T.m[1][b] + FBS[a][2] * T.m[2][b];
      break;
    }
    case VIEKF_DEPTH: {
      int i = global_to_local_feature_id(id);
      double rho = x[xZ + i * 5 + 4];
      zh[0] = 1.0 / rho;
      H(0, dxZ + 3 * i + 2) = -1.0 / (rho * rho);
      break;
    }
    case VIEKF_INV_DEPTH: {
      int i = global_to_local_feature_id(id);
      zh[0] = x[xZ + i * 5 + 4];
      H(0, dxZ + 3 * i + 2) = 1.0;
      break;
    }
    default:  // PIXEL_VEL: empty stub in the reference (vi_ekf_meas.cpp:388-395)
      break;
  }
}

// ---- vi_ekf_meas.cpp:196-278 ------------------------------------------------------------------------------
int VIEKF::update(Meas& meas) {
  meas.handled = true;
  // A feature-indexed model whose id is no longer tracked (a queued measurement replayed after keep_only_features dropped
  // its feature) makes the reference index the state with -1 (UB, vi_ekf_meas.cpp:337,356,371,381).  Defined here, as in
  // the engine, as MEAS_INVALID with no state change.
  if ((meas.type == VIEKF_FEAT || meas.type == VIEKF_QZETA || meas.type == VIEKF_DEPTH || meas.type == VIEKF_INV_DEPTH) &&
      global_to_local_feature_id(meas.id) < 0)
    return VIEKF_MEAS_INVALID;
  for (int i = 0; i < 4; i++) zhat_[i] = 0;
  H_.zero();
  K_.zero();
  h(meas.type, x_[i_], zhat_, H_, meas.id);

  double residual[4] = {0, 0, 0, 0};
  if (meas.type == VIEKF_QZETA) {
    q_feat_boxminus(Quat(meas.z), Quat(zhat_), residual);
  } else if (meas.type == VIEKF_ATT) {
    V3 d = Quat(meas.z).boxminus(Quat(zhat_));
    for (int i = 0; i < 3; i++) residual[i] = d[i];
  } else {
    for (int i = 0; i < meas.zdim; i++) residual[i] = meas.z[i] - zhat_[i];
  }
  const int r = meas.rdim, n = MAX_DX;
  Mat& P = P_[i_];

  if (meas.active) {
    // innov = (H P H^T + R)^-1
    Mat HP(r, n);
    for (int j = 0; j < n; j++)
      for (int a = 0; a < r; a++) {
        double s = 0;
        for (int k = 0; k < n; k++) s += H_(a, k) * P(k, j);
        HP(a, j) = s;
      }
    double S[9], Si[9];
    for (int a = 0; a < r; a++)
      for (int b = 0; b < r; b++) {
        double s = 0;
        for (int k = 0; k < n; k++) s += HP(a, k) * H_(b, k);
        S[a + b * r] = s + meas.R[a + b * 3];
      }
    small_inverse(S, r, Si);
    double mahal = 0;
    {
      double tmp[3];
      for (int b = 0; b < r; b++) {
        double s = 0;
        for (int a = 0; a < r; a++) s += residual[a] * Si[a + b * r];
        tmp[b] = s;
      }
      for (int b = 0; b < r; b++) mahal += tmp[b] * residual[b];
    }
    min_gate_margin = std::min(min_gate_margin, std::fabs(mahal - 9.0));
    if (mahal > 9.0) {
      n_gated++;
      return VIEKF_MEAS_GATED;  // early return: no fix_depth
    }
    // K = P H^T innov
    Mat PHt(n, r);
    for (int a = 0; a < r; a++)
      for (int i = 0; i < n; i++) {
        double s = 0;
        for (int k = 0; k < n; k++) s += P(i, k) * H_(a, k);
        PHt(i, a) = s;
      }
    bool nans = false;
    for (int b = 0; b < r; b++)
      for (int i = 0; i < n; i++) {
        double s = 0;
        for (int a = 0; a < r; a++) s += PHt(i, a) * Si[a + b * r];
        K_(i, b) = s;
        if (s != s) nans = true;
      }
    for (double v : H_.d)
      if (v != v) nans = true;
    if (!nans) {
      std::vector<double> d(n, 0.0);
      for (int i = 0; i < n; i++) {
        double s = 0;
        for (int a = 0; a < r; a++) s += K_(i, a) * residual[a];
        d[i] = cfg_.use_partial_update ? lambda_[i] * s : s;
      }
      boxplus(x_[i_], d, xp_);
      x_[i_] = xp_;
      // A_ = I - K H
      for (int j = 0; j < n; j++)
        for (int i = 0; i < n; i++) {
          double s = 0;
          for (int a = 0; a < r; a++) s += K_(i, a) * H_(a, j);
          A_(i, j) = (i == j ? 1.0 : 0.0) - s;
        }
      gemm(A_, P, T1_);
      gemm_nt(T1_, A_, T3_);
      // K R K^T
      Mat KR(n, r);
      for (int b = 0; b < r; b++)
        for (int i = 0; i < n; i++) {
          double s = 0;
          for (int a = 0; a < r; a++) s += K_(i, a) * meas.R[a + b * 3];
          KR(i, b) = s;
        }
      for (int j = 0; j < n; j++)
        for (int i = 0; i < n; i++) {
          double s = 0;
          for (int a = 0; a < r; a++) s += KR(i, a) * K_(j, a);
          T2_(i, j) = s;
        }
      if (cfg_.use_partial_update) {
        for (size_t k = 0; k < P.d.size(); k++) P.d[k] += Lambda_.d[k] * ((T3_.d[k] + T2_.d[k]) - P.d[k]);
      } else {
        for (size_t k = 0; k < P.d.size(); k++) P.d[k] = T3_.d[k] + T2_.d[k];
      }
    } else {
      flags |= VIEKF_FLAG_NAN_GAIN;
      n_nan_gain++;
    }
  }
  fix_depth();
  return VIEKF_MEAS_SUCCESS;
}

// ---- vi_ekf_meas.cpp:130-194 --------------------------------------------------------------------------------
int VIEKF::add_measurement(double t, const double* z, int zdim, int type, const double* R, int rdim, bool active, int id,
                           double depth) {
  if (t < start_t_) return VIEKF_MEAS_INVALID;
  for (int i = 0; i < zdim; i++)
    if (z[i] != z[i]) return VIEKF_MEAS_NAN;
  if (type == VIEKF_FEAT && id >= 0) {
    if (std::find(current_feature_ids_.begin(), current_feature_ids_.end(), id) == current_feature_ids_.end()) {
      init_feature(z, id, depth);
      return VIEKF_MEAS_NEW_FEATURE;
    }
  }
  auto z_it = zbuf_.begin();
  while (z_it != zbuf_.end()) {
    if (z_it->t < t) break;
    z_it++;
  }
  Meas m;
  std::memset(&m, 0, sizeof(m));
  m.t = t;
  m.type = type;
  m.zdim = zdim;
  m.rdim = rdim;
  for (int a = 0; a < rdim; a++)
    for (int b = 0; b < rdim; b++) m.R[a + b * 3] = R[a + b * rdim];
  for (int i = 0; i < zdim; i++) m.z[i] = z[i];
  m.active = active;
  m.id = id;
  m.depth = depth;
  m.handled = false;
  if (z_it == zbuf_.begin())
    zbuf_.push_front(m);
  else
    zbuf_.insert(z_it, m);
  return VIEKF_MEAS_SUCCESS;
}

// ---- vi_ekf_meas.cpp:6-127 ------------------------------------------------------------------------------------
void VIEKF::handle_measurements(std::vector<int>* gated) {
  if (gated) gated->clear();
  if (zbuf_.size() == 0) return;
  auto z_it = zbuf_.end() - 1;
  while (z_it->handled == true && z_it != zbuf_.begin()) z_it--;
  if (z_it == zbuf_.begin() && z_it->handled) return;
  if (z_it->t > u_.begin()->first) return;  // measurement from the future

  auto u_it = u_.begin();
  while (u_it != u_.end()) {
    if (z_it->t > u_it->first) break;
    u_it++;
  }
  if (u_it == u_.end() || z_it->t <= u_it->first) return;  // not enough input history (reference derefs end(): UB)

  int i = HIST;
  while (i > 0) {
    if (t_[(i_ + i) % HIST] <= u_it->first) {
      i_ = (i_ + i) % HIST;
      break;
    }
    i--;
  }
  if (i == 0) {
    zbuf_.erase(z_it);
    return;
  }
  // Replay (vi_ekf_meas.cpp:73-117).  The stored inputs are already in the body frame and propagate_state
  // rotates them by q_b_u again -- reproduced literally (quirk Q17 in DESIGN.md).
  u_it--;
  while (u_it != u_.begin()) {
    while (z_it->t <= u_it->first) {
      if (t_[i_] < z_it->t) propagate_state(u_it->second.data(), z_it->t, false);
      if (!z_it->handled) {
        int result = update(*z_it);
        if (gated != nullptr && result == VIEKF_MEAS_GATED && z_it->type == VIEKF_FEAT) gated->push_back(z_it->id);
      }
      if (z_it != zbuf_.begin()) {
        z_it--;
        while (u_it->first < z_it->t && u_it != u_.begin()) {
          propagate_state(u_it->second.data(), u_it->first, false);
          u_it--;
        }
      } else {
        while (u_it != u_.begin()) {
          propagate_state(u_it->second.data(), u_it->first, false);
          u_it--;
        }
        break;
      }
    }
  }
  propagate_state(u_it->second.data(), u_it->first, false);
  while ((int)zbuf_.size() > LEN_MEAS_HIST) zbuf_.pop_back();
  while ((int)u_.size() > LEN_STATE_HIST) u_.pop_back();
}

// ---- vi_ekf_feat.cpp:6-47 ---------------------------------------------------------------------------------------
bool VIEKF::init_feature(const double l[2], int id, double depth) {
  (void)id;  // ignored by the reference: ids are assigned internally
  if (len_features_ >= NF) return false;
  double lc0 = l[0] - cfg_.cam_center[0], lc1 = l[1] - cfg_.cam_center[1];
  V3 zeta = v3(lc0, lc1 * (cam_F_[1] / cam_F_[0]), cam_F_[0]);  // Q4: literal
  zeta = zeta / norm(zeta);
  Quat qzeta = Quat::from_two_unit_vectors(v3(0, 0, 1), zeta);
  double init_depth = depth;
  if (depth != depth) init_depth = 2.0 * cfg_.min_depth;
  current_feature_ids_.push_back(next_feature_id_);
  next_feature_id_ += 1;
  len_features_ += 1;
  std::vector<double>& x = x_[i_];
  Mat& P = P_[i_];
  int x_max = xZ + 5 * len_features_;
  for (int k = 0; k < 4; k++) x[x_max - 5 + k] = qzeta.q[k];
  x[x_max - 1] = 1.0 / init_depth;
  int dx_max = dxZ + 3 * len_features_;
  for (int i = dx_max - 3; i < dx_max; i++)
    for (int j = 0; j < dx_max - 3; j++) {
      P(i, j) = 0.0;
      P(j, i) = 0.0;
    }
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) P(dx_max - 3 + i, dx_max - 3 + j) = (i == j) ? P0_feat_[i] : 0.0;
  return true;
}

// ---- vi_ekf_feat.cpp:50-73 ---------------------------------------------------------------------------------------
void VIEKF::clear_feature(int id) {
  int local = global_to_local_feature_id(id);
  if (local < 0) return;  // the reference would index with -1 (UB); callers only pass tracked ids
  int xZETA_i = xZ + 5 * local, dxZETA_i = dxZ + 3 * local;
  current_feature_ids_.erase(current_feature_ids_.begin() + local);
  len_features_ -= 1;
  int dx_max = dxZ + 3 * len_features_;
  std::vector<double>& x = x_[i_];
  Mat& P = P_[i_];
  const int n = MAX_DX;
  if (local < len_features_) {
    for (int i = xZETA_i; i + 5 < MAX_X; i++) x[i] = x[i + 5];
    for (int j = 0; j < n; j++)
      for (int i = dxZETA_i; i + 3 < n; i++) P(i, j) = P(i + 3, j);
    for (int j = dxZETA_i; j + 3 < n; j++)
      for (int i = 0; i < n; i++) P(i, j) = P(i, j + 3);
  }
  for (int i = xZ + 5 * len_features_; i < MAX_X; i++) x[i] = 0.0;
  for (int j = dx_max; j < n; j++)
    for (int i = 0; i < n; i++) P(i, j) = 0.0;
  for (int j = 0; j < n; j++)
    for (int i = dx_max; i < n; i++) P(i, j) = 0.0;
}

// ---- vi_ekf_feat.cpp:81-142 --------------------------------------------------------------------------------------
void VIEKF::keep_only_features(const std::vector<int>& features) {
  std::vector<int> to_remove;
  int overlap = 0;
  for (size_t local = 0; local < current_feature_ids_.size(); local++) {
    bool keep = false;
    for (size_t i = 0; i < features.size(); i++) {
      if (current_feature_ids_[local] == features[i]) {
        keep = true;
        if (cfg_.use_keyframe_reset) {
          for (size_t j = 0; j < keyframe_features_.size(); j++)
            if (keyframe_features_[j] == features[i]) {
              overlap++;
              break;
            }
        }
        break;
      }
    }
    if (!keep) to_remove.push_back(current_feature_ids_[local]);
  }
  for (int id : to_remove) clear_feature(id);
  if (cfg_.use_keyframe_reset && keyframe_features_.size() > 0 &&
      (double)overlap / (double)keyframe_features_.size() < cfg_.keyframe_overlap_threshold) {
    keyframe_reset();
    keyframe_features_ = features;
  } else if (cfg_.use_keyframe_reset && keyframe_features_.size() == 0) {
    keyframe_features_ = features;
  }
}

// ---- vi_ekf_kfr.cpp -------------------------------------------------------------------------------------------
// Adj of the current node pose (xform.h:226-235): [[R, [t]x R],[0, R]]; P_new = P_prev + Adj^T cov Adj.
void VIEKF::propagate_global_covariance(const double* P_prev, const Edge& e, double* P_new) const {
  double Adj[36] = {0};
  M3 R = node_q_.R();
  M3 SR = mul(skew(v3(node_t_[0], node_t_[1], node_t_[2])), R);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      Adj[i + 6 * j] = R.m[i][j];
      Adj[i + 6 * (j + 3)] = SR.m[i][j];
      Adj[(i + 3) + 6 * (j + 3)] = R.m[i][j];
    }
  double T[36];
  for (int i = 0; i < 6; i++)
    for (int j = 0; j < 6; j++) {
      double s = 0;
      for (int k = 0; k < 6; k++) s += Adj[k + 6 * i] * e.cov[k + 6 * j];
      T[i + 6 * j] = s;
    }
  double out[36];
  for (int i = 0; i < 6; i++)
    for (int j = 0; j < 6; j++) {
      double s = 0;
      for (int k = 0; k < 6; k++) s += T[i + 6 * k] * Adj[k + 6 * j];
      out[i + 6 * j] = P_prev[i + 6 * j] + s;
    }
  for (int i = 0; i < 36; i++) P_new[i] = out[i];
}

void VIEKF::keyframe_reset() {
  std::vector<double>& x = x_[i_];
  Mat& P = P_[i_];
  Edge edge;
  std::memset(&edge, 0, sizeof(edge));
  for (int i = 0; i < 3; i++) edge.t[i] = x[xPOS + i];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) edge.cov[i + 6 * j] = P(xPOS + i, xPOS + j);
  for (int i = 0; i < 3; i++) x[xPOS + i] = 0.0;
  Quat q_i2b(&x[xATT]);
  double yaw = q_i2b.yaw(), roll = q_i2b.roll(), pitch = q_i2b.pitch();
  Quat qe = Quat::from_euler(0, 0, yaw);
  for (int i = 0; i < 4; i++) edge.q[i] = qe.q[i];
  edge.cov[5 + 6 * 5] = P(xATT + 2, xATT + 2);  // literal: indexes P with the *state* index xATT+2 = 8 (vi_ekf_kfr.cpp:126)
  Quat qn = Quat::from_euler(roll, pitch, 0.0);
  for (int i = 0; i < 4; i++) x[xATT + i] = qn.q[i];
  double cp = std::cos(roll), sp = std::sin(roll), tt = std::tan(pitch);
  const int n = MAX_DX;
  A_.zero();
  for (int i = 0; i < n; i++) A_(i, i) = 1.0;
  for (int i = 0; i < 3; i++) A_(dxPOS + i, dxPOS + i) = 0.0;
  double Nb[3][3] = {{1, sp * tt, cp * tt}, {0, cp * cp, -cp * sp}, {0, -cp * sp, sp * sp}};
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) A_(dxATT + i, dxATT + j) = Nb[i][j];
  gemm(A_, P, T1_);
  gemm_nt(T1_, A_, T3_);
  P.d = T3_.d;
  propagate_global_covariance(global_cov_, edge, global_cov_);
  // current_node_global_pose_ = current_node_global_pose_ * edge.transform  (xform.h:236-244)
  V3 et = v3(edge.t[0], edge.t[1], edge.t[2]);
  V3 nt = v3(node_t_[0], node_t_[1], node_t_[2]) + node_q_.rota(et);
  for (int i = 0; i < 3; i++) node_t_[i] = nt[i];
  node_q_ = node_q_.otimes(qe);
}

void VIEKF::keyframe_reset(const std::vector<double>& xm, std::vector<double>& xp, Mat& N) {
  x_[i_] = xm;
  keyframe_reset();
  xp = x_[i_];
  N = A_;
}

void VIEKF::get_global_pose(double t[3], double q[4]) const {
  const std::vector<double>& x = x_[i_];
  V3 rt = v3(x[xPOS], x[xPOS + 1], x[xPOS + 2]);
  V3 nt = v3(node_t_[0], node_t_[1], node_t_[2]) + node_q_.rota(rt);
  Quat qq = node_q_.otimes(Quat(&x[xATT]));
  for (int i = 0; i < 3; i++) t[i] = nt[i];
  for (int i = 0; i < 4; i++) q[i] = qq.q[i];
}

void VIEKF::get_global_cov(double cov[36]) const {
  const Mat& P = P_[i_];
  Edge rel;
  std::memset(&rel, 0, sizeof(rel));
  // literal: the reference indexes P with state indices xPOS / xATT = 0 / 6 (vi_ekf_kfr.cpp:30-34)
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      rel.cov[i + 6 * j] = P(xPOS + i, xPOS + j);
      rel.cov[(i + 3) + 6 * j] = P(xATT + i, xPOS + j);
      rel.cov[i + 6 * (j + 3)] = P(xPOS + i, xATT + j);
      rel.cov[(i + 3) + 6 * (j + 3)] = P(xATT + i, xATT + j);
    }
  propagate_global_covariance(global_cov_, rel, cov);
}

// ---- in-order tape semantics (include/viekf_batch.h) ---------------------------------------------------------------
int VIEKF::update_direct(int type, const double* z, int zdim, const double* R, int rdim, bool active, int id, double depth) {
  // add_measurement's bookkeeping (vi_ekf_meas.cpp:133-147) followed immediately by update()
  for (int i = 0; i < zdim; i++)
    if (z[i] != z[i]) return VIEKF_MEAS_NAN;
  if (type == VIEKF_FEAT && id >= 0 && global_to_local_feature_id(id) < 0) {
    init_feature(z, id, depth);
    return VIEKF_MEAS_NEW_FEATURE;
  }
  // A feature-indexed model with an untracked id makes the reference index the state with -1 (UB,
  // vi_ekf_meas.cpp:337,356,371,381); the tape semantics define it as MEAS_INVALID with no state change.
  if ((type == VIEKF_FEAT || type == VIEKF_QZETA || type == VIEKF_DEPTH || type == VIEKF_INV_DEPTH) && global_to_local_feature_id(id) < 0)
    return VIEKF_MEAS_INVALID;
  Meas m;
  std::memset(&m, 0, sizeof(m));
  m.t = t_[i_];
  m.type = type;
  m.zdim = zdim;
  m.rdim = rdim;
  for (int a = 0; a < rdim; a++)
    for (int b = 0; b < rdim; b++) m.R[a + b * 3] = R[a + b * rdim];
  for (int i = 0; i < zdim; i++) m.z[i] = z[i];
  m.active = active;
  m.id = id;
  m.depth = depth;
  return update(m);
}

void VIEKF::perturb_x0(uint64_t seed, uint64_t ginst) {
  std::vector<double> d(MAX_DX, 0.0), out(MAX_X, 0.0);
  for (int k = 0; k < 8; k++) {
    double nn[2];
    viekf_normal_pair(seed, ginst, VIEKF_NOISE_EVENT_X0, k, nn);
    d[2 * k] = std::sqrt(cfg_.P0[2 * k]) * nn[0];
    d[2 * k + 1] = std::sqrt(cfg_.P0[2 * k + 1]) * nn[1];
  }
  const int keep = len_features_;
  len_features_ = 0;
  out = x_[i_];
  boxplus(x_[i_], d, out);
  len_features_ = keep;
  x_[i_] = out;
}

void VIEKF::run_tape(const viekf_event* ev, int64_t n, const int32_t* id_pool, const double* values, uint8_t* results,
                     double* stats, int n_buckets, uint64_t seed, const double* bias_range, uint64_t ginst, int64_t ev_first) {
  double bias[6] = {0, 0, 0, 0, 0, 0};
  if (seed && bias_range)
    for (int k = 0; k < 6; k++) bias[k] = bias_range[k] * viekf_uniform_pm1(seed, ginst, VIEKF_NOISE_EVENT_BIAS, k);
  for (int64_t e = ev_first; e < ev_first + n; e++) {
    const viekf_event& E = ev[e];
    const double* zsrc = (E.noise >= 0 && values) ? values + 8 * (int64_t)E.noise : E.z;
    // per-instance value resolution (include/viekf_batch.h: viekf_event.sigma, viekf_tape.seed / imu_bias_range)
    double zbuf[8];
    for (int k = 0; k < 8; k++) zbuf[k] = zsrc[k];
    if (seed && (E.kind == VIEKF_EV_PROP || E.kind == VIEKF_EV_UPD || E.kind == VIEKF_EV_INIT)) {
      for (int k = 0; k < 8; k += 2)
        if (E.sigma[k] != 0.0 || E.sigma[k + 1] != 0.0) {
          double nn[2];
          const uint32_t key = (E.kind == VIEKF_EV_PROP && E.aux >= 0) ? (0x80000000u | (uint32_t)E.aux) : (uint32_t)e;
          viekf_normal_pair(seed, ginst, key, (uint32_t)(k >> 1), nn);
          zbuf[k] += E.sigma[k] * nn[0];
          zbuf[k + 1] += E.sigma[k + 1] * nn[1];
        }
      if (E.kind == VIEKF_EV_PROP)
        for (int k = 0; k < 6; k++)
          if (bias_range && bias_range[k] != 0.0) zbuf[k] += bias[k];
    }
    const double* zv = zbuf;
    int res = VIEKF_MEAS_NONE;
    switch (E.kind) {
      case VIEKF_EV_PROP:
        propagate_state(zv, E.t, false);
        break;
      case VIEKF_EV_UPD: {
        double z[4] = {0, 0, 0, 0};
        for (int i = 0; i < E.zdim; i++) z[i] = (E.aux == 1) ? ub_last_[i] : zv[i];
        res = update_direct(E.type, z, E.zdim, E.R, E.rdim, E.active != 0, E.id, E.z[4]);
        break;
      }
      case VIEKF_EV_INIT:
        res = init_feature(zv, E.id, E.z[4]) ? 1 : 0;
        break;
      case VIEKF_EV_CLEAR:
        clear_feature(E.id);
        break;
      case VIEKF_EV_KEEP: {
        std::vector<int> ids(id_pool + E.ids_off, id_pool + E.ids_off + E.n_ids);
        keep_only_features(ids);
        break;
      }
      case VIEKF_EV_KFRESET:
        keyframe_reset();
        break;
      case VIEKF_EV_STAT: {
        if (!stats || E.id < 0 || E.id >= n_buckets) break;
        double* s = stats + (size_t)E.id * VIEKF_STAT_WIDTH;
        // a filter with a NaN in its live state or covariance is flagged, counted, and kept out of every sum
        if (check_nan_state()) {
          s[38] += 1.0;
          break;
        }
        double xt[17];
        for (int i = 0; i < 8; i++) xt[i] = E.z[i];
        for (int i = 0; i < 9; i++) xt[8 + i] = E.R[i];
        if (E.noise >= 0 && values)
          for (int i = 0; i < 6; i++) xt[xB_A + i] = zv[i];
        if (seed) {  // the instance's generated IMU bias, in the body frame
          V3 ba = q_b_u_.rota(v3(bias[0], bias[1], bias[2])), bg = q_b_u_.rota(v3(bias[3], bias[4], bias[5]));
          for (int i = 0; i < 3; i++) {
            xt[xB_A + i] += ba[i];
            xt[xB_G + i] += bg[i];
          }
        }
        // The truth is global; after a keyframe reset the filter's position and attitude are relative to the current node
        // (vi_ekf_kfr.cpp:56-157).  The truth is therefore expressed in the filter's node frame -- the inverse of the
        // get_global_pose composition (vi_ekf_kfr.cpp:14-22): p_rel = node_q.rotp(p - node_t), q_rel = node_q^-1 * q --
        // which is the identity map until the first reset.
        {
          V3 pr = node_q_.rotp(v3(xt[0] - node_t_[0], xt[1] - node_t_[1], xt[2] - node_t_[2]));
          Quat qr = node_q_.inverse().otimes(Quat(&xt[xATT]));
          for (int i = 0; i < 3; i++) xt[i] = pr[i];
          for (int i = 0; i < 4; i++) xt[xATT + i] = qr.q[i];
        }
        const std::vector<double>& x = x_[i_];
        double er[16];
        for (int i = 0; i < 6; i++) er[i] = xt[i] - x[i];
        V3 d = Quat(&xt[xATT]).boxminus(Quat(&x[xATT]));
        for (int i = 0; i < 3; i++) er[6 + i] = d[i];
        for (int i = 0; i < 7; i++) er[9 + i] = xt[xB_A + i] - x[xB_A + i];
        s[0] += 1.0;
        for (int i = 0; i < 16; i++) {
          s[1 + i] += er[i];
          s[17 + i] += er[i] * er[i];
        }
        // NEES via Cholesky of the base block (16) and of the pose block (pos+att, 6)
        auto nees = [&](const int* idx, int m) {
          // NEES is defined only for a positive-definite block: a non-positive pivot (the position block is exactly
          // zero right after a keyframe reset) makes it NaN
          double L[16 * 16], y[16];
          const Mat& P = P_[i_];
          double q = 0;
          for (int i = 0; i < m; i++) {
            for (int j = 0; j <= i; j++) {
              double sum = P(idx[i], idx[j]);
              for (int k = 0; k < j; k++) sum -= L[i * 16 + k] * L[j * 16 + k];
              if (i == j) {
                if (!(sum > 0.0)) return (double)NAN;
                L[i * 16 + j] = std::sqrt(sum);
              } else {
                L[i * 16 + j] = sum / L[j * 16 + j];
              }
            }
            double sum = er[idx[i]];
            for (int k = 0; k < i; k++) sum -= L[i * 16 + k] * y[k];
            y[i] = sum / L[i * 16 + i];
            q += y[i] * y[i];
          }
          return q;
        };
        int all[16], pose[6] = {0, 1, 2, 6, 7, 8};
        for (int i = 0; i < 16; i++) all[i] = i;
        // an undefined (non-finite) NEES is left out of the sum; [37] / [39] count the defined ones
        const double q16 = nees(all, 16), q6 = nees(pose, 6);
        if (std::isfinite(q16)) {
          s[33] += q16;
          s[37] += 1.0;
        }
        if (std::isfinite(q6)) {
          s[34] += q6;
          s[39] += 1.0;
        }
        s[35] += (double)n_gated;
        s[36] += (double)n_fix;
        break;
      }
      default:
        break;
    }
    if (results) results[e] = (uint8_t)res;
  }
  check_nan_state();  // end of the slice: the engine's executors latch VIEKF_FLAG_NAN_STATE here too
}

}  // namespace viekf_oracle
