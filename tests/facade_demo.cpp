// Drives viekf::VIEKF (include/viekf_facade.hpp) with the call pattern of the reference's test/vi_ekf_test.cpp:24-33,58:
// imu callback -> propagate_state, feature callback -> add_measurement(FEAT), plus keep_only_features as the ROS image
// callback does (src/vi_ekf_ros.cpp:282).  Prints the final state; exit code 3 = no CUDA device (the engine has no CPU path).
#include <cstdio>
#include <cstring>
#include "viekf_facade.hpp"

// `facade_demo queue`: the streaming pattern of src/vi_ekf_ros.cpp:174-198,282-309 with the delayed-measurement queue on
// (frames stamped 12 ms in the past, handle_measurements after every IMU sample).
static int run_queue() {
  viekf_config cfg = viekf::default_config();
  cfg.use_keyframe_reset = 0;
  viekf::VIEKF ekf(cfg, 4);
  ekf.enable_queue();
  const double R2[4] = {0.25, 0, 0, 0.25};
  const double Racc[4] = {1, 0, 0, 1};
  int n_gated = 0, n_new = 0;
  std::vector<int> gated;
  for (int k = 0; k <= 40; k++) {
    const double t = 0.005 * (k + 1);
    const double u[6] = {0.1, -0.05, -9.80665, 0.01, 0.02, -0.01};
    ekf.propagate_state(u, t);
    const double acc[2] = {u[0], u[1]};
    if (k > 0) ekf.add_measurement(t, acc, 2, viekf::ACC, Racc, 2, true);
    if (k % 8 == 0)
      for (int id = 0; id < 3; id++) {
        const double pix[2] = {300.0 + 20 * id + 0.1 * k, 220.0 + 10 * id};
        n_new += (ekf.add_measurement(k == 0 ? t : t - 0.012, pix, 2, viekf::FEAT, R2, 2, true, id, 4.0) == viekf::MEAS_NEW_FEATURE);
      }
    ekf.handle_measurements(&gated);
    n_gated += (int)gated.size();
  }
  std::vector<double> x = ekf.get_state();
  std::printf("len %d new %d ok 0 gated %d t %.3f\n", ekf.get_len_features(), n_new, n_gated, ekf.get_time());
  std::printf("x");
  for (size_t i = 0; i < x.size(); i++) std::printf(" %.17g", x[i]);
  std::printf("\n");
  return 0;
}

int main(int argc, char** argv) {
  try {
    if (argc > 1 && std::strcmp(argv[1], "queue") == 0) return run_queue();
    viekf_config cfg = viekf::default_config();
    viekf::VIEKF ekf(cfg, 4);
    const double R2[4] = {0.25, 0, 0, 0.25};
    const double Racc[4] = {1, 0, 0, 1};
    int n_new = 0, n_ok = 0, n_gated = 0;
    for (int k = 0; k <= 40; k++) {
      const double t = 0.005 * k;
      const double u[6] = {0.1, -0.05, -9.80665, 0.01, 0.02, -0.01};
      ekf.propagate_state(u, t);
      const double acc[2] = {u[0], u[1]};
      ekf.add_measurement(t, acc, 2, viekf::ACC, Racc, 2, true);
      if (k % 8 == 4) {
        for (int id = 0; id < 3; id++) {
          const double pix[2] = {300.0 + 20 * id, 220.0 + 10 * id};
          viekf::meas_result_t r = ekf.add_measurement(t, pix, 2, viekf::FEAT, R2, 2, true, id, 4.0);
          n_new += (r == viekf::MEAS_NEW_FEATURE);
          n_ok += (r == viekf::MEAS_SUCCESS);
          n_gated += (r == viekf::MEAS_GATED);
        }
        ekf.keep_only_features({0, 1, 2});
      }
    }
    ekf.handle_measurements();
    std::vector<double> x = ekf.get_state();
    std::vector<int> ids = ekf.tracked_features();
    std::printf("len %d new %d ok %d gated %d t %.3f\n", ekf.get_len_features(), n_new, n_ok, n_gated, ekf.get_time());
    std::printf("x");
    for (size_t i = 0; i < x.size(); i++) std::printf(" %.17g", x[i]);
    std::printf("\n");
    return 0;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return std::strstr(e.what(), "no CUDA device") ? 3 : 1;
  }
}
