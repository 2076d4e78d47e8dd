// viekf_oracle.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement (C++17, no Eigen) of the filter core of byu-magicc/VI-EKF, used only as the parity
// checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
// The product path (vi-ekf_b200/csrc) never includes, links or calls anything in this directory.
//
// PARITY STATUS: "parity unpinned" for x / P trajectories -- the reference holds no golden vector for a
// propagate or an update, it cannot be compiled in this image (Eigen3, yaml-cpp, gtest are
// absent) and its dense arithmetic lives in un-vendored, version-unpinned Eigen3 (CMakeLists.txt:17).  What IS
// pinned: the reference's own property tests re-run against this restatement (tests/test_oracle_properties.py:
// manifold identities, finite-difference checks of A, G and every H, keyframe-reset Jacobian -- reference
// test/jac_test.cpp:245-487) and the three known-answer geometry cases of lib/geometry/src/test.cpp:45-73,154-181.
//
// The restatement performs the same dense MAX_DX x MAX_DX arithmetic, in the same operation order, with the
// same branches, thresholds and quirks as the reference files cited at each function.  Matrices are
// column-major like Eigen's default.  Dense products accumulate over k in ascending order (Eigen's internal
// order is unspecified, so bit-equality with an Eigen build was never attainable; contract is rel 1e-9).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <vector>
#include <algorithm>
#include <array>
#include "../include/viekf_batch.h"

namespace viekf_oracle {

// ------------------------------------------------------------------------------------------------
// small fixed-size helpers (3-vectors, 3x3 row-major arrays m[r][c])
// ------------------------------------------------------------------------------------------------
struct V3 { double v[3]; double& operator[](int i) { return v[i]; } double operator[](int i) const { return v[i]; } };
struct M3 { double m[3][3]; };

inline V3 v3(double a, double b, double c) { return V3{{a, b, c}}; }
inline V3 operator+(const V3& a, const V3& b) { return v3(a[0] + b[0], a[1] + b[1], a[2] + b[2]); }
inline V3 operator-(const V3& a, const V3& b) { return v3(a[0] - b[0], a[1] - b[1], a[2] - b[2]); }
inline V3 operator*(double s, const V3& a) { return v3(s * a[0], s * a[1], s * a[2]); }
inline V3 operator*(const V3& a, double s) { return v3(a[0] * s, a[1] * s, a[2] * s); }
inline V3 operator/(const V3& a, double s) { return v3(a[0] / s, a[1] / s, a[2] / s); }
inline double dot(const V3& a, const V3& b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline V3 cross(const V3& a, const V3& b) {
  return v3(a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]);
}
inline double norm(const V3& a) { return std::sqrt(dot(a, a)); }

// skew: reference lib/geometry/include/geometry/support.h:58-68
inline M3 skew(const V3& v) {
  M3 s = {{{0.0, -v[2], v[1]}, {v[2], 0.0, -v[0]}, {-v[1], v[0], 0.0}}};
  return s;
}
inline M3 mul(const M3& a, const M3& b) {
  M3 c;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      double s = 0.0;
      for (int k = 0; k < 3; k++) s += a.m[i][k] * b.m[k][j];
      c.m[i][j] = s;
    }
  return c;
}
inline M3 transpose(const M3& a) {
  M3 c;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) c.m[i][j] = a.m[j][i];
  return c;
}
inline M3 scale(const M3& a, double s) {
  M3 c;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) c.m[i][j] = a.m[i][j] * s;
  return c;
}
inline M3 add(const M3& a, const M3& b) {
  M3 c;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) c.m[i][j] = a.m[i][j] + b.m[i][j];
  return c;
}
inline M3 sub(const M3& a, const M3& b) {
  M3 c;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) c.m[i][j] = a.m[i][j] - b.m[i][j];
  return c;
}
inline V3 mul(const M3& a, const V3& v) {
  return v3(a.m[0][0] * v[0] + a.m[0][1] * v[1] + a.m[0][2] * v[2], a.m[1][0] * v[0] + a.m[1][1] * v[1] + a.m[1][2] * v[2],
            a.m[2][0] * v[0] + a.m[2][1] * v[1] + a.m[2][2] * v[2]);
}
inline M3 eye3() {
  M3 s = {{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}};
  return s;
}

// ------------------------------------------------------------------------------------------------
// Quat: reference lib/geometry/include/geometry/quat.h.  Storage [w,x,y,z] (quat.h:75-78).
// ------------------------------------------------------------------------------------------------
struct Quat {
  double q[4];
  Quat() : q{1.0, 0.0, 0.0, 0.0} {}
  Quat(double w, double x, double y, double z) : q{w, x, y, z} {}
  explicit Quat(const double* p) : q{p[0], p[1], p[2], p[3]} {}
  double w() const { return q[0]; }
  double x() const { return q[1]; }
  double y() const { return q[2]; }
  double z() const { return q[3]; }
  V3 bar() const { return v3(q[1], q[2], q[3]); }

  // quat.h:376-391 (rota = R^T v), :362-374 (rotp = R v)
  V3 rota(const V3& v) const {
    V3 t = 2.0 * cross(v, bar());
    return v - w() * t + cross(t, bar());
  }
  V3 rotp(const V3& v) const {
    V3 t = 2.0 * cross(v, bar());
    return v + w() * t + cross(t, bar());
  }
  // quat.h:306-322, passive rotation matrix
  M3 R() const {
    double wx = w() * x(), wy = w() * y(), wz = w() * z();
    double xx = x() * x(), xy = x() * y(), xz = x() * z();
    double yy = y() * y(), yz = y() * z(), zz = z() * z();
    M3 o = {{{1. - 2. * yy - 2. * zz, 2. * xy + 2. * wz, 2. * xz - 2. * wy},
             {2. * xy - 2. * wz, 1. - 2. * xx - 2. * zz, 2. * yz + 2. * wx},
             {2. * xz + 2. * wy, 2. * yz - 2. * wx, 1. - 2. * xx - 2. * yy}}};
    return o;
  }
  // quat.h:398-406
  Quat inverse() const { return Quat(q[0], -q[1], -q[2], -q[3]); }
  // quat.h:419-428 (Hamilton product)
  Quat otimes(const Quat& o) const {
    return Quat(w() * o.w() - x() * o.x() - y() * o.y() - z() * o.z(), w() * o.x() + x() * o.w() + y() * o.z() - z() * o.y(),
                w() * o.y() - x() * o.z() + y() * o.w() + z() * o.x(), w() * o.z() + x() * o.y() - y() * o.x() + z() * o.w());
  }
  // quat.h:128-144
  static Quat exp(const V3& v) {
    double n = norm(v);
    Quat o;
    if (n > 1e-4) {
      double s = std::sin(n / 2.0) / n;
      o = Quat(std::cos(n / 2.0), s * v[0], s * v[1], s * v[2]);
    } else {
      o = Quat(1.0, v[0] / 2.0, v[1] / 2.0, v[2] / 2.0);
      double nn = std::sqrt(o.q[0] * o.q[0] + o.q[1] * o.q[1] + o.q[2] * o.q[2] + o.q[3] * o.q[3]);
      for (int i = 0; i < 4; i++) o.q[i] /= nn;
    }
    return o;
  }
  // quat.h:146-162
  static V3 log(const Quat& p) {
    V3 v = p.bar();
    double w = p.w();
    double n = norm(v);
    if (n < 1e-8) return v3(0, 0, 0);
    return (2.0 * std::atan2(n, w)) * v / n;
  }
  // quat.h:430-434
  Quat boxplus(const V3& d) const { return otimes(exp(d)); }
  // quat.h:436-445  (this [-] o)
  V3 boxminus(const Quat& o) const {
    Quat dq = o.inverse().otimes(*this);
    if (dq.w() < 0.0)
      for (int i = 0; i < 4; i++) dq.q[i] *= -1.0;
    return log(dq);
  }
  // quat.h:233-255
  static Quat from_two_unit_vectors(const V3& u, const V3& v) {
    double d = dot(u, v);
    if (d < 0.99999999 && d > -0.99999999) {
      double invs = 1.0 / std::sqrt(2.0 * (1.0 + d));
      V3 xyz = cross(u, v * invs);
      Quat o(0.5 / invs, xyz[0], xyz[1], xyz[2]);
      double nn = std::sqrt(o.q[0] * o.q[0] + o.q[1] * o.q[1] + o.q[2] * o.q[2] + o.q[3] * o.q[3]);
      for (int i = 0; i < 4; i++) o.q[i] /= nn;
      return o;
    } else if (d < -0.99999999) {
      return Quat(0, 1, 0, 0);
    }
    return Quat(1, 0, 0, 0);
  }
  // quat.h:217-231
  static Quat from_euler(double roll, double pitch, double yaw) {
    double cp = std::cos(roll / 2.0), ct = std::cos(pitch / 2.0), cs = std::cos(yaw / 2.0);
    double sp = std::sin(roll / 2.0), st = std::sin(pitch / 2.0), ss = std::sin(yaw / 2.0);
    return Quat(cp * ct * cs + sp * st * ss, sp * ct * cs - cp * st * ss, cp * st * cs + sp * ct * ss, cp * ct * ss - sp * st * cs);
  }
  // quat.h:162-195 (from_R)
  static Quat from_R(const M3& M) {
    const double(*m)[3] = M.m;
    double tr = m[0][0] + m[1][1] + m[2][2];
    if (tr > 0) {
      double S = std::sqrt(tr + 1.0) * 2.;
      return Quat(0.25 * S, (m[1][2] - m[2][1]) / S, (m[2][0] - m[0][2]) / S, (m[0][1] - m[1][0]) / S);
    } else if ((m[0][0] > m[1][1]) && (m[0][0] > m[2][2])) {
      double S = std::sqrt(1.0 + m[0][0] - m[1][1] - m[2][2]) * 2.;
      return Quat((m[1][2] - m[2][1]) / S, 0.25 * S, (m[1][0] + m[0][1]) / S, (m[2][0] + m[0][2]) / S);
    } else if (m[1][1] > m[2][2]) {
      double S = std::sqrt(1.0 + m[1][1] - m[0][0] - m[2][2]) * 2.;
      return Quat((m[2][0] - m[0][2]) / S, (m[1][0] + m[0][1]) / S, 0.25 * S, (m[2][1] + m[1][2]) / S);
    }
    double S = std::sqrt(1.0 + m[2][2] - m[0][0] - m[1][1]) * 2.;
    return Quat((m[0][1] - m[1][0]) / S, (m[2][0] + m[0][2]) / S, (m[2][1] + m[1][2]) / S, 0.25 * S);
  }
  // quat.h:267-290
  double roll() const { return std::atan2(2.0 * (w() * x() + y() * z()), 1.0 - 2.0 * (x() * x() + y() * y())); }
  double pitch() const {
    double val = 2.0 * (w() * y() - x() * z());
    if (std::fabs(val) > 1.0) return std::copysign(1.0, val) * M_PI / 2.0;
    return std::asin(val);
  }
  double yaw() const { return std::atan2(2.0 * (w() * z() + x() * y()), 1.0 - 2.0 * (y() * y() + z() * z())); }
};

// reference include/math_helper.h:14-22 ; T_zeta = [q.rota(e_x) q.rota(e_y)] via doublerota (quat.h:336-346)
inline V3 zeta_of(const Quat& q) { return q.rota(v3(0, 0, 1)); }
struct M32 { double m[3][2]; };
inline M32 T_zeta(const Quat& q) {
  M32 T;
  V3 a = q.rota(v3(1, 0, 0)), b = q.rota(v3(0, 1, 0));
  for (int i = 0; i < 3; i++) { T.m[i][0] = a[i]; T.m[i][1] = b[i]; }
  return T;
}
// math_helper.h:25-43  (q_j [-] q_i)
inline void q_feat_boxminus(const Quat& qj, const Quat& qi, double out[2]) {
  V3 zi = zeta_of(qi), zj = zeta_of(qj);
  if (norm(zi - zj) > 1e-8) {
    V3 s = cross(zi, zj);
    s = s / norm(s);
    double theta = std::acos(dot(zi, zj));
    M32 T = T_zeta(qi);
    V3 ts = theta * s;
    out[0] = T.m[0][0] * ts[0] + T.m[1][0] * ts[1] + T.m[2][0] * ts[2];
    out[1] = T.m[0][1] * ts[0] + T.m[1][1] * ts[1] + T.m[2][1] * ts[2];
  } else {
    out[0] = out[1] = 0.0;
  }
}
// math_helper.h:45-48
inline Quat q_feat_boxplus(const Quat& q, const double dq[2]) {
  M32 T = T_zeta(q);
  V3 d = v3(T.m[0][0] * dq[0] + T.m[0][1] * dq[1], T.m[1][0] * dq[0] + T.m[1][1] * dq[1], T.m[2][0] * dq[0] + T.m[2][1] * dq[1]);
  return Quat::exp(d).otimes(q);
}

// ------------------------------------------------------------------------------------------------
// dense column-major matrix
// ------------------------------------------------------------------------------------------------
struct Mat {
  int r = 0, c = 0;
  std::vector<double> d;
  Mat() {}
  Mat(int r_, int c_) : r(r_), c(c_), d((size_t)r_ * c_, 0.0) {}
  double& operator()(int i, int j) { return d[(size_t)i + (size_t)j * r]; }
  double operator()(int i, int j) const { return d[(size_t)i + (size_t)j * r]; }
  void zero() { std::fill(d.begin(), d.end(), 0.0); }
};

struct Meas {  // measurement_t, reference include/vi_ekf.h:167-179
  double t;
  int type;
  double z[4];
  double R[9];  // 3x3 column-major, top-left rdim x rdim used
  bool active;
  int id;
  double depth;
  int zdim, rdim;
  bool handled;
};

struct Edge {  // edge_t (include/vi_ekf.h:127-130): transform = (t, q), cov 6x6 column-major
  double t[3];
  double q[4];
  double cov[36];
};

static const int LEN_STATE_HIST = 250;  // include/vi_ekf.h:50
static const int LEN_MEAS_HIST = 200;   // include/vi_ekf.h:51

class VIEKF {
 public:
  // NF = NUM_FEATURES (compile-time in the reference, include/vi_ekf.h:39-48).  hist = LEN_STATE_HIST; the CPU
  // baseline uses a short ring for in-order tapes (the ring depth does not change in-order results).
  VIEKF(const viekf_config& cfg, int NF, int hist = LEN_STATE_HIST);

  const int NF, MAX_X, MAX_DX, HIST;

  // --- reference public surface (include/vi_ekf.h:243-337) ---
  void propagate_state(const double u[6], double t, bool save_input = true);  // vi_ekf.cpp:262-318
  void dynamics(const std::vector<double>& x, const double u[6], bool state = true, bool jac = true);  // vi_ekf_dyn.cpp:14-135
  void boxplus(const std::vector<double>& x, const std::vector<double>& dx, std::vector<double>& out) const;  // vi_ekf_helper.cpp:88-98
  void boxminus(const std::vector<double>& x1, const std::vector<double>& x2, std::vector<double>& out) const;  // :100-111
  int add_measurement(double t, const double* z, int zdim, int type, const double* R, int rdim, bool active = false, int id = -1,
                      double depth = NAN);                     // vi_ekf_meas.cpp:130-194
  void handle_measurements(std::vector<int>* gated = nullptr);  // vi_ekf_meas.cpp:6-127
  int update(Meas& m);                                         // vi_ekf_meas.cpp:196-278
  void h(int type, const std::vector<double>& x, double zhat[4], Mat& H, int id) const;  // vi_ekf_meas.cpp:281-395
  bool init_feature(const double l[2], int id, double depth = -1.0);  // vi_ekf_feat.cpp:6-47
  void clear_feature(int id);                                        // vi_ekf_feat.cpp:50-73
  void keep_only_features(const std::vector<int>& features);         // vi_ekf_feat.cpp:81-142
  void keyframe_reset();                                             // vi_ekf_kfr.cpp:56-157
  void keyframe_reset(const std::vector<double>& xm, std::vector<double>& xp, Mat& N);  // vi_ekf_kfr.cpp:6-12
  void fix_depth();                                                  // vi_ekf_helper.cpp:128-156
  int global_to_local_feature_id(int id) const;                      // vi_ekf_helper.cpp:114-125
  bool nans_in_the_house() const;                                    // vi_ekf_error.cpp:6-19
  bool check_nan_state();                                            // latches VIEKF_FLAG_NAN_STATE; returns the sticky flag
  void get_global_pose(double t[3], double q[4]) const;              // vi_ekf_kfr.cpp:14-22
  void get_global_cov(double cov[36]) const;                         // vi_ekf_kfr.cpp:24-37

  const std::vector<double>& get_state() const { return x_[i_]; }
  const Mat& get_covariance() const { return P_[i_]; }
  std::vector<double>& state_mut() { return x_[i_]; }
  Mat& cov_mut() { return P_[i_]; }
  int get_len_features() const { return len_features_; }
  const std::vector<int>& tracked_features() const { return current_feature_ids_; }
  double time() const { return t_[i_]; }
  void set_x0(const double* x0) { for (int i = 0; i < 17; i++) x_[i_][i] = x0[i]; }

  // In-order tape execution (semantics of include/viekf_batch.h).  values = this instance's value stream.
  // results[e] receives meas_result_t or VIEKF_MEAS_NONE.  stats may be null.
  // seed / bias_range / ginst: generated per-instance noise (include/viekf_noise.h); ev_first: absolute index of ev[0]'s
  // successor range [ev_first, ev_first + n) inside the tape (noise is keyed by the absolute event index).
  void run_tape(const viekf_event* ev, int64_t n, const int32_t* id_pool, const double* values, uint8_t* results, double* stats,
                int n_buckets, uint64_t seed = 0, const double* bias_range = nullptr, uint64_t ginst = 0, int64_t ev_first = 0);
  void perturb_x0(uint64_t seed, uint64_t ginst);  // x <- x [+] sqrt(P0) o n   (viekf_batch_perturb_x0)
  int update_direct(int type, const double* z, int zdim, const double* R, int rdim, bool active, int id, double depth);

  // workspace, public so that the ported jac_test properties can read them (the reference's dynamics()
  // overload copies them out, vi_ekf_dyn.cpp:5-11)
  Mat A_, G_;
  std::vector<double> dx_;
  Mat H_, K_;
  double zhat_[4];
  double ub_last_[6] = {0, 0, 0, 0, 0, 0};  // body-frame IMU input of the last propagate_state call
  uint32_t flags = 0;
  uint32_t n_gated = 0, n_fix = 0, n_nan_gain = 0;
  double min_gate_margin = 1e300;  // min |mahal - 9| seen (SURVEY 7 "hard parts": reject seeds with tiny margins)
  double min_depth_margin = 1e300; // min distance of rho to the fix_depth thresholds 0 and 100

 private:
  viekf_config cfg_;
  int i_ = 0;
  std::vector<std::vector<double>> x_;
  std::vector<Mat> P_;
  std::vector<double> t_;
  std::deque<std::pair<double, std::array<double, 6>>> u_;
  std::deque<Meas> zbuf_;
  Mat Qx_;
  double Qu_[6];
  std::vector<double> lambda_;
  Mat Lambda_;
  double P0_feat_[3];
  double start_t_ = NAN;
  int len_features_ = 0;
  int next_feature_id_ = 0;
  std::vector<int> current_feature_ids_;
  std::vector<int> keyframe_features_;
  std::vector<double> xp_;
  Mat T1_, T2_, T3_;  // dense temporaries
  Quat q_b_c_, q_b_u_;
  V3 p_b_c_;
  double cam_F_[2];  // cam_F_(0,0), cam_F_(1,1)
  // global pose bookkeeping (vi_ekf_kfr.cpp)
  double node_t_[3] = {0, 0, 0};
  Quat node_q_;
  double global_cov_[36];
  void propagate_global_covariance(const double* P_prev, const Edge& e, double* P_new) const;  // vi_ekf_kfr.cpp:49-55
};

}  // namespace viekf_oracle
