// oracle_capi.cpp -- TEST INFRASTRUCTURE.  ctypes-friendly C entry points over viekf_oracle::VIEKF.
// Loaded only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
#include "viekf_oracle.hpp"
#include <atomic>
#include <chrono>
#include <thread>

using viekf_oracle::VIEKF;
using viekf_oracle::Quat;
using viekf_oracle::V3;
using viekf_oracle::M3;
using viekf_oracle::Mat;

extern "C" {

void* vo_create(const viekf_config* cfg, int NF, int hist) { return new VIEKF(*cfg, NF, hist); }
void vo_destroy(void* h) { delete (VIEKF*)h; }
void vo_propagate(void* h, const double* u, double t, int save) { ((VIEKF*)h)->propagate_state(u, t, save != 0); }
int vo_add_measurement(void* h, double t, const double* z, int zdim, int type, const double* R, int rdim, int active, int id,
                       double depth) {
  return ((VIEKF*)h)->add_measurement(t, z, zdim, type, R, rdim, active != 0, id, depth);
}
int vo_handle_measurements(void* h, int* gated, int cap) {
  std::vector<int> g;
  ((VIEKF*)h)->handle_measurements(&g);
  for (int i = 0; i < (int)g.size() && i < cap; i++) gated[i] = g[i];
  return (int)g.size();
}
int vo_update_direct(void* h, int type, const double* z, int zdim, const double* R, int rdim, int active, int id, double depth) {
  return ((VIEKF*)h)->update_direct(type, z, zdim, R, rdim, active != 0, id, depth);
}
int vo_init_feature(void* h, const double* l, int id, double depth) { return ((VIEKF*)h)->init_feature(l, id, depth) ? 1 : 0; }
void vo_clear_feature(void* h, int id) { ((VIEKF*)h)->clear_feature(id); }
void vo_keep_only_features(void* h, const int* ids, int n) { ((VIEKF*)h)->keep_only_features(std::vector<int>(ids, ids + n)); }
void vo_keyframe_reset(void* h) { ((VIEKF*)h)->keyframe_reset(); }
void vo_fix_depth(void* h) { ((VIEKF*)h)->fix_depth(); }
void vo_get_state(void* h, double* x) {
  VIEKF* f = (VIEKF*)h;
  std::memcpy(x, f->get_state().data(), sizeof(double) * f->MAX_X);
}
void vo_set_state(void* h, const double* x) {
  VIEKF* f = (VIEKF*)h;
  std::memcpy(f->state_mut().data(), x, sizeof(double) * f->MAX_X);
}
void vo_get_cov(void* h, double* P) {
  VIEKF* f = (VIEKF*)h;
  std::memcpy(P, f->get_covariance().d.data(), sizeof(double) * f->MAX_DX * f->MAX_DX);
}
void vo_set_cov(void* h, const double* P) {
  VIEKF* f = (VIEKF*)h;
  std::memcpy(f->cov_mut().d.data(), P, sizeof(double) * f->MAX_DX * f->MAX_DX);
}
int vo_len_features(void* h) { return ((VIEKF*)h)->get_len_features(); }
void vo_tracked_features(void* h, int* ids) {
  const std::vector<int>& v = ((VIEKF*)h)->tracked_features();
  for (size_t i = 0; i < v.size(); i++) ids[i] = v[i];
}
double vo_time(void* h) { return ((VIEKF*)h)->time(); }
unsigned vo_flags(void* h) {
  ((VIEKF*)h)->check_nan_state();
  return ((VIEKF*)h)->flags;
}
void vo_counters(void* h, unsigned* c) {
  VIEKF* f = (VIEKF*)h;
  c[0] = f->n_gated;
  c[1] = f->n_fix;
  c[2] = f->n_nan_gain;
}
void vo_margins(void* h, double* m) {
  VIEKF* f = (VIEKF*)h;
  m[0] = f->min_gate_margin;
  m[1] = f->min_depth_margin;
}
void vo_dynamics(void* h, const double* x, const double* u, double* dx, double* A, double* G) {
  VIEKF* f = (VIEKF*)h;
  std::vector<double> xv(x, x + f->MAX_X);
  f->dynamics(xv, u, true, true);
  std::memcpy(dx, f->dx_.data(), sizeof(double) * f->MAX_DX);
  std::memcpy(A, f->A_.d.data(), sizeof(double) * f->MAX_DX * f->MAX_DX);
  std::memcpy(G, f->G_.d.data(), sizeof(double) * f->MAX_DX * 6);
}
void vo_boxplus(void* h, const double* x, const double* dx, double* out) {
  VIEKF* f = (VIEKF*)h;
  std::vector<double> xv(x, x + f->MAX_X), d(dx, dx + f->MAX_DX), o(f->MAX_X, 0.0);
  f->boxplus(xv, d, o);
  std::memcpy(out, o.data(), sizeof(double) * f->MAX_X);
}
void vo_boxminus(void* h, const double* x1, const double* x2, double* out) {
  VIEKF* f = (VIEKF*)h;
  std::vector<double> a(x1, x1 + f->MAX_X), b(x2, x2 + f->MAX_X), o(f->MAX_DX, 0.0);
  f->boxminus(a, b, o);
  std::memcpy(out, o.data(), sizeof(double) * f->MAX_DX);
}
void vo_h(void* h, int type, const double* x, int id, double* zhat, double* H) {
  VIEKF* f = (VIEKF*)h;
  std::vector<double> xv(x, x + f->MAX_X);
  Mat Hm(3, f->MAX_DX);
  double zh[4] = {0, 0, 0, 0};
  f->h(type, xv, zh, Hm, id);
  std::memcpy(zhat, zh, sizeof(zh));
  std::memcpy(H, Hm.d.data(), sizeof(double) * 3 * f->MAX_DX);
}
void vo_keyframe_reset_jac(void* h, const double* xm, double* xp, double* N) {
  VIEKF* f = (VIEKF*)h;
  std::vector<double> a(xm, xm + f->MAX_X), o(f->MAX_X, 0.0);
  Mat Nm(f->MAX_DX, f->MAX_DX);
  f->keyframe_reset(a, o, Nm);
  std::memcpy(xp, o.data(), sizeof(double) * f->MAX_X);
  std::memcpy(N, Nm.d.data(), sizeof(double) * f->MAX_DX * f->MAX_DX);
}
void vo_global_pose(void* h, double* t3q4, double* cov36) {
  VIEKF* f = (VIEKF*)h;
  f->get_global_pose(t3q4, t3q4 + 3);
  f->get_global_cov(cov36);
}
void vo_run_tape(void* h, const viekf_event* ev, int64_t n, const int32_t* id_pool, const double* values, uint8_t* results,
                 double* stats, int n_buckets, uint64_t seed, const double* bias_range, uint64_t ginst, int64_t ev_first) {
  ((VIEKF*)h)->run_tape(ev, n, id_pool, values, results, stats, n_buckets, seed, bias_range, ginst, ev_first);
}
void vo_perturb_x0(void* h, uint64_t seed, uint64_t ginst) { ((VIEKF*)h)->perturb_x0(seed, ginst); }

// CPU baseline: `count` instances of the same tape, one instance per thread at a time over `nthreads` host
// threads.  Returns wall seconds.  x_out ([MAX_X x count]) / P_out ([MAX_DX^2 x count]) / results_out /
// stats (n_buckets x VIEKF_STAT_WIDTH, summed over instances) may be null.
double vo_run_batch(const viekf_config* cfg, int NF, const viekf_event* ev, int64_t n_ev, const int32_t* id_pool,
                    const double* values, int64_t values_stride, const double* x0 /*[17 x count] or null*/, int64_t count,
                    int nthreads, double* x_out, double* P_out, uint8_t* results_out, double* stats, int n_buckets,
                    int* len_out, int* ids_out, unsigned* flags_out, double* margins_out, uint64_t seed, const double* bias_range,
                    int64_t inst_off, uint64_t perturb_seed) {
  std::atomic<int64_t> next(0);
  std::vector<std::vector<double>> tstats(nthreads, std::vector<double>((size_t)std::max(n_buckets, 0) * VIEKF_STAT_WIDTH, 0.0));
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&](int tid) {
    for (;;) {
      int64_t b = next.fetch_add(1);
      if (b >= count) break;
      VIEKF f(*cfg, NF, 2);
      if (x0) f.set_x0(x0 + 17 * b);
      if (perturb_seed) f.perturb_x0(perturb_seed, (uint64_t)(inst_off + b));
      f.run_tape(ev, n_ev, id_pool, values ? values + b * values_stride : nullptr, results_out ? results_out + b * n_ev : nullptr,
                 n_buckets > 0 ? tstats[tid].data() : nullptr, n_buckets, seed, bias_range, (uint64_t)(inst_off + b), 0);
      if (x_out) std::memcpy(x_out + (size_t)b * f.MAX_X, f.get_state().data(), sizeof(double) * f.MAX_X);
      if (P_out) std::memcpy(P_out + (size_t)b * f.MAX_DX * f.MAX_DX, f.get_covariance().d.data(), sizeof(double) * f.MAX_DX * f.MAX_DX);
      if (len_out) len_out[b] = f.get_len_features();
      if (ids_out) {
        for (int i = 0; i < NF; i++) ids_out[(size_t)b * NF + i] = -1;
        for (size_t i = 0; i < f.tracked_features().size(); i++) ids_out[(size_t)b * NF + i] = f.tracked_features()[i];
      }
      if (flags_out) {
        flags_out[4 * b] = f.flags;
        flags_out[4 * b + 1] = f.n_gated;
        flags_out[4 * b + 2] = f.n_fix;
        flags_out[4 * b + 3] = f.n_nan_gain;
      }
      if (margins_out) {
        margins_out[2 * b] = f.min_gate_margin;
        margins_out[2 * b + 1] = f.min_depth_margin;
      }
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < nthreads; i++) th.emplace_back(work, i);
  work(0);
  for (auto& t : th) t.join();
  double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (stats && n_buckets > 0)
    for (int t = 0; t < nthreads; t++)
      for (size_t i = 0; i < tstats[t].size(); i++) stats[i] += tstats[t][i];
  return secs;
}

// geometry library probes for the known-answer tests (lib/geometry/src/test.cpp)
//  op: 0 rota(a=q,b=v)->3  1 rotp->3  2 R(q)->9 row-major  3 exp(b=v)->4  4 log(q)->3  5 otimes(a,b)->4
//      6 boxplus(q,v)->4  7 boxminus(q1,q2)->3  8 from_two_unit_vectors(a=u3,b=v3)->4  9 from_euler(a=rpy)->4
//      10 euler(q)->3  11 from_R(a=9 row-major)->4  12 inverse->4  13 q_feat_boxplus(q, b=dq2)->4
//      14 q_feat_boxminus(qj=a, qi=b)->2  15 T_zeta(q)->6 row-major 3x2  16 zeta(q)->3  17 from_axis_angle(a=axis, b[0]=angle)->4
void vo_quat_op(int op, const double* a, const double* b, double* out) {
  auto V = [](const double* p) { return viekf_oracle::v3(p[0], p[1], p[2]); };
  auto put3 = [&](const V3& v) { for (int i = 0; i < 3; i++) out[i] = v[i]; };
  auto put4 = [&](const Quat& q) { for (int i = 0; i < 4; i++) out[i] = q.q[i]; };
  switch (op) {
    case 0: put3(Quat(a).rota(V(b))); break;
    case 1: put3(Quat(a).rotp(V(b))); break;
    case 2: { M3 R = Quat(a).R(); for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) out[3 * i + j] = R.m[i][j]; break; }
    case 3: put4(Quat::exp(V(b))); break;
    case 4: put3(Quat::log(Quat(a))); break;
    case 5: put4(Quat(a).otimes(Quat(b))); break;
    case 6: put4(Quat(a).boxplus(V(b))); break;
    case 7: put3(Quat(a).boxminus(Quat(b))); break;
    case 8: put4(Quat::from_two_unit_vectors(V(a), V(b))); break;
    case 9: put4(Quat::from_euler(a[0], a[1], a[2])); break;
    case 10: { Quat q(a); out[0] = q.roll(); out[1] = q.pitch(); out[2] = q.yaw(); break; }
    case 11: { M3 R; for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) R.m[i][j] = a[3 * i + j]; put4(Quat::from_R(R)); break; }
    case 12: put4(Quat(a).inverse()); break;
    case 13: put4(viekf_oracle::q_feat_boxplus(Quat(a), b)); break;
    case 14: viekf_oracle::q_feat_boxminus(Quat(a), Quat(b), out); break;
    case 15: { auto T = viekf_oracle::T_zeta(Quat(a)); for (int i = 0; i < 3; i++) for (int j = 0; j < 2; j++) out[2 * i + j] = T.m[i][j]; break; }
    case 16: put3(viekf_oracle::zeta_of(Quat(a))); break;
    case 17: {  // quat.h:197-208
      double a2 = b[0] / 2.0, s = std::sin(a2);
      Quat q(std::cos(a2), a[0] * s, a[1] * s, a[2] * s);
      double n = std::sqrt(q.q[0] * q.q[0] + q.q[1] * q.q[1] + q.q[2] * q.q[2] + q.q[3] * q.q[3]);
      for (int i = 0; i < 4; i++) out[i] = q.q[i] / n;
      break;
    }
  }
}

int vo_sizeof_event() { return (int)sizeof(viekf_event); }
int vo_sizeof_config() { return (int)sizeof(viekf_config); }
}
