// viekf_facade.hpp -- header-only C++ facade with the reference's method names over the C ABI of viekf_batch.h.
//
// `viekf::VIEKF` is ONE filter (a batch of B = 1) behind the public surface of the reference's `vi_ekf::VIEKF`
// (reference include/vi_ekf.h:243-337), so a call site of the reference -- test/vi_ekf_test.cpp:24-33,58,
// src/vi_ekf_ros.cpp:174-198,282-309 -- compiles against it with the Eigen types replaced by plain arrays; when Eigen is
// installed (__has_include(<Eigen/Core>)) the Eigen-typed overloads of the reference are provided as well.
// `viekf::BatchVIEKF` is the same surface over B Monte-Carlo copies (inputs broadcast or one column per copy).
//
// Semantics preserved from the reference: the first propagate_state only latches the time (vi_ekf.cpp:274-279); feature
// ids are assigned internally (vi_ekf_feat.cpp:29-31); a FEAT measurement with an untracked id initialises the feature and
// returns MEAS_NEW_FEATURE (vi_ekf_meas.cpp:140-147).  Two modes: by default measurements are applied in order at the filter's
// current time and handle_measurements() has nothing left to do; after enable_queue() the reference's
// delayed-measurement queue is in force (vi_ekf_meas.cpp:6-194): add_measurement queues by time stamp, handle_measurements
// rewinds the history ring and replays, exactly as the reference's streaming callers expect (src/vi_ekf_ros.cpp:174-198).
// Errors: the reference prints to cerr or exits; here every failing ABI call throws std::runtime_error(viekf_last_error()).
#pragma once
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <array>
#include <vector>
#include "viekf_batch.h"
#if defined(__has_include)
#if __has_include(<Eigen/Core>)
#include <Eigen/Core>
#define VIEKF_FACADE_HAS_EIGEN 1
#endif
#endif

namespace viekf {

// reference include/vi_ekf.h:132-138 (same numeric values)
enum meas_result_t { MEAS_SUCCESS = VIEKF_MEAS_SUCCESS, MEAS_GATED = VIEKF_MEAS_GATED, MEAS_NAN = VIEKF_MEAS_NAN,
                     MEAS_INVALID = VIEKF_MEAS_INVALID, MEAS_NEW_FEATURE = VIEKF_MEAS_NEW_FEATURE };
// reference include/vi_ekf.h:167-179
enum measurement_type_t { ACC = VIEKF_ACC, ALT = VIEKF_ALT, ATT = VIEKF_ATT, POS = VIEKF_POS, VEL = VIEKF_VEL, QZETA = VIEKF_QZETA,
                          FEAT = VIEKF_FEAT, PIXEL_VEL = VIEKF_PIXEL_VEL, DEPTH = VIEKF_DEPTH, INV_DEPTH = VIEKF_INV_DEPTH };

inline viekf_config default_config() {
  viekf_config c;
  viekf_config_default(&c);
  return c;
}

class BatchVIEKF {
 public:
  // reference: VIEKF::init(...) / load(param_file) (vi_ekf.cpp:18-155); `cfg` carries what load() reads from params/ekf.yaml
  BatchVIEKF(const viekf_config& cfg, int64_t B, int num_features, int device = 0) : B_(B), N_(num_features) {
    check(viekf_batch_create(&cfg, B, num_features, device, &h_));
  }
  ~BatchVIEKF() {
    if (h_) viekf_batch_destroy(h_);
  }
  BatchVIEKF(const BatchVIEKF&) = delete;
  BatchVIEKF& operator=(const BatchVIEKF&) = delete;

  int64_t size() const { return B_; }
  int num_features() const { return N_; }
  int max_x() const { return 17 + 5 * N_; }    // reference xZ + 5 NUM_FEATURES
  int max_dx() const { return 16 + 3 * N_; }   // reference dxZ + 3 NUM_FEATURES
  viekf_batch* handle() { return h_; }

  // vi_ekf.cpp:157-184
  void set_x0(const double* x0, bool broadcast = true) { check(viekf_batch_set_x0(h_, x0, broadcast ? 1 : 0)); }
  // vi_ekf.cpp:262-318: u = [acc(3), gyro(3)] in the IMU frame, one column shared by all copies or [6 x B]
  void propagate_state(const double* u, double t, bool broadcast = true, bool save_input = true) {
    if (queue_)
      check(viekf_batch_queue_propagate(h_, u, t, save_input ? 1 : 0));  // one shared IMU sample
    else
      check(viekf_batch_propagate(h_, u, broadcast ? 1 : 0, t));
  }
  // include/vi_ekf.h:50-51: LEN_STATE_HIST / LEN_MEAS_HIST; must precede the first propagate_state
  void enable_queue(int len_state_hist = 250, int len_meas_hist = 200) {
    check(viekf_batch_queue_enable(h_, len_state_hist, len_meas_hist));
    queue_ = true;
  }
  bool queue_enabled() const { return queue_; }
  // vi_ekf_meas.cpp:130-194 + 196-278; returns one meas_result_t per copy
  std::vector<uint8_t> add_measurement(double t, const double* z, int zdim, measurement_type_t type, const double* R, int rdim,
                                       bool active = false, int id = -1, double depth = NAN, bool broadcast = true) {
    std::vector<uint8_t> res((size_t)B_);
    if (queue_) {
      int32_t r = 0;
      check(viekf_batch_add_measurement(h_, t, (int32_t)type, z, zdim, R, rdim, active ? 1 : 0, id, depth, &r));
      res.assign((size_t)B_, (uint8_t)r);
      return res;
    }
    check(viekf_batch_update(h_, (int32_t)type, z, zdim, broadcast ? 1 : 0, R, rdim, active ? 1 : 0, id, depth, res.data()));
    return res;
  }
  // vi_ekf_meas.cpp:6-127; gated_feature_ids (may be null) receives the FEAT ids gated for copy `instance`
  void handle_measurements(std::vector<int>* gated_feature_ids = nullptr, int64_t instance = 0) {
    if (gated_feature_ids) gated_feature_ids->clear();
    if (!queue_) return;  // in-order application: nothing is queued
    const int cap = 512;
    std::vector<int32_t> ty(cap), ids(cap);
    std::vector<uint8_t> res((size_t)cap * (size_t)B_);
    int32_t n = 0;
    check(viekf_batch_handle_measurements(h_, ty.data(), ids.data(), res.data(), cap, &n));
    if (gated_feature_ids)
      for (int k = 0; k < n && k < cap; k++)
        if (ty[k] == VIEKF_FEAT && res[(size_t)k * B_ + instance] == VIEKF_MEAS_GATED) gated_feature_ids->push_back(ids[k]);
  }
  // vi_ekf_feat.cpp:6-142
  std::vector<uint8_t> init_feature(const double* pixel, int id, double depth = NAN, bool broadcast = true) {
    std::vector<uint8_t> ok((size_t)B_);
    check(viekf_batch_init_feature(h_, pixel, broadcast ? 1 : 0, id, depth, ok.data()));
    return ok;
  }
  void clear_feature(int id) { check(viekf_batch_clear_feature(h_, id)); }
  void keep_only_features(const std::vector<int>& ids) {
    std::vector<int32_t> v(ids.begin(), ids.end());
    check(viekf_batch_keep_only_features(h_, v.data(), (int32_t)v.size()));
  }
  // vi_ekf_kfr.cpp:56-157
  void keyframe_reset() { check(viekf_batch_keyframe_reset(h_)); }
  // read-back (vi_ekf.h:272-286): x is [max_x x count] column per copy, P is count dense column-major max_dx^2 blocks
  std::vector<double> get_state(int64_t first = 0, int64_t count = -1) {
    if (count < 0) count = B_ - first;
    std::vector<double> x((size_t)(count * max_x()));
    check(viekf_batch_get_state(h_, first, count, x.data()));
    return x;
  }
  std::vector<double> get_covariance(int64_t first = 0, int64_t count = -1) {
    if (count < 0) count = B_ - first;
    std::vector<double> P((size_t)(count * max_dx() * max_dx()));
    check(viekf_batch_get_covariance(h_, first, count, P.data()));
    return P;
  }
  std::vector<int> tracked_features(int64_t instance = 0) {
    int32_t len = 0;
    std::vector<int32_t> ids((size_t)(N_ > 0 ? N_ : 1));
    check(viekf_batch_get_features(h_, instance, 1, &len, ids.data()));
    return std::vector<int>(ids.begin(), ids.begin() + len);
  }
  int get_len_features(int64_t instance = 0) { return (int)tracked_features(instance).size(); }
  // vi_ekf.cpp:202-260 (feature slots beyond len_features are NaN; an untracked id gives NaN)
  std::vector<double> get_covariance_diagonal(int64_t first = 0, int64_t count = -1) {
    if (count < 0) count = B_ - first;
    std::vector<double> d((size_t)(count * max_dx()));
    check(viekf_batch_get_covariance_diagonal(h_, first, count, d.data()));
    return d;
  }
  std::vector<double> get_depths(int64_t first = 0, int64_t count = -1) {
    if (count < 0) count = B_ - first;
    std::vector<double> d((size_t)(count * (N_ > 0 ? N_ : 1)), NAN);
    check(viekf_batch_get_depths(h_, first, count, d.data()));
    return d;
  }
  std::vector<double> get_zetas(int64_t first = 0, int64_t count = -1) {
    if (count < 0) count = B_ - first;
    std::vector<double> z((size_t)(3 * count * (N_ > 0 ? N_ : 1)), NAN);
    check(viekf_batch_get_zetas(h_, first, count, z.data()));
    return z;
  }
  std::vector<double> get_qzetas(int64_t first = 0, int64_t count = -1) {
    if (count < 0) count = B_ - first;
    std::vector<double> q((size_t)(4 * count * (N_ > 0 ? N_ : 1)), NAN);
    check(viekf_batch_get_qzetas(h_, first, count, q.data()));
    return q;
  }
  double get_depth(int id, int64_t instance = 0) {
    double d = NAN;
    check(viekf_batch_get_depth(h_, id, instance, 1, &d));
    return d;
  }
  std::array<double, 2> get_feat(int id, int64_t instance = 0) {
    std::array<double, 2> px{NAN, NAN};
    check(viekf_batch_get_feat(h_, id, instance, 1, px.data()));
    return px;
  }
  void set_imu_bias(const double (&b_g)[3], const double (&b_a)[3]) { check(viekf_batch_set_imu_bias(h_, b_g, b_a)); }
  void set_drag_term(bool use_drag_term) { check(viekf_batch_set_drag_term(h_, use_drag_term ? 1 : 0)); }
  bool get_drag_term() const { return viekf_batch_get_drag_term(h_) != 0; }
  // VIEKF(param_file) (vi_ekf.cpp:101-155)
  static viekf_config load(const std::string& param_file) {
    viekf_config c;
    check(viekf_config_load_yaml(param_file.c_str(), &c));
    return c;
  }
  double get_time(int64_t instance = 0) {
    double t = 0;
    check(viekf_batch_get_time(h_, instance, 1, &t));
    return t;
  }

 protected:
  static void check(int rc) {
    if (rc != 0) throw std::runtime_error(std::string("viekf: ") + viekf_last_error());
  }
  viekf_batch* h_ = nullptr;
  bool queue_ = false;
  int64_t B_;
  int N_;
};

// One filter with the reference's call signatures (arrays in place of Eigen types).
class VIEKF : public BatchVIEKF {
 public:
  explicit VIEKF(const viekf_config& cfg, int num_features = 12 /* reference default NUM_FEATURES */, int device = 0)
      : BatchVIEKF(cfg, 1, num_features, device) {}
  void propagate_state(const double (&u)[6], double t, bool save_input = true) { BatchVIEKF::propagate_state(u, t, true, save_input); }
  meas_result_t add_measurement(double t, const double* z, int zdim, measurement_type_t type, const double* R, int rdim, bool active = false,
                                int id = -1, double depth = NAN) {
    return (meas_result_t)BatchVIEKF::add_measurement(t, z, zdim, type, R, rdim, active, id, depth, true)[0];
  }
  bool init_feature(const double (&l)[2], int id, double depth = NAN) { return BatchVIEKF::init_feature(l, id, depth, true)[0] != 0; }
#ifdef VIEKF_FACADE_HAS_EIGEN
  // the reference's own signatures (include/vi_ekf.h:294-309)
  void propagate_state(const Eigen::Matrix<double, 6, 1>& u, double t, bool save_input = true) {
    BatchVIEKF::propagate_state(u.data(), t, true, save_input);
  }
  meas_result_t add_measurement(double t, const Eigen::VectorXd& z, measurement_type_t type, const Eigen::MatrixXd& R, bool active = false,
                                int id = -1, double depth = NAN) {
    return add_measurement(t, z.data(), (int)z.rows(), type, R.data(), (int)R.rows(), active, id, depth);
  }
  bool init_feature(const Eigen::Vector2d& l, int id, double depth = NAN) { return BatchVIEKF::init_feature(l.data(), id, depth, true)[0] != 0; }
  Eigen::VectorXd get_state_eigen() {
    std::vector<double> x = get_state();
    return Eigen::Map<Eigen::VectorXd>(x.data(), (Eigen::Index)x.size());
  }
#endif
};

}  // namespace viekf
