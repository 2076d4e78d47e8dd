// viekf_capi.cu -- the C ABI of include/viekf_batch.h over the tape-executor kernels.
//
// Every entry point names the reference method it stands in for in include/viekf_batch.h.  There is no CPU
// fallback: without a CUDA device viekf_batch_create fails with VIEKF_ERR_NODEVICE.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <array>
#include <deque>
#include <mutex>
#include <string>
#include <vector>
#include "viekf_engine.h"

using vk::FilterMeta;
using vk::KParams;

static thread_local std::string g_err;

struct viekf_dev_tape {
  viekf_event* ev = nullptr;
  int64_t n_events = 0;
  int32_t* id_pool = nullptr;
  double* values = nullptr;
  int64_t vstride = 0;
  uint64_t seed = 0;
  double bias_range[6] = {0, 0, 0, 0, 0, 0};
};

// Delayed-measurement queue of the streaming call pattern (vi_ekf_meas.cpp:6-194): the (x, P) history ring lives in HBM
// (the handle's state pointers kp.x / kp.P always address slot i), the time ring and the two deques on the host.
struct QueuedMeas {
  double t, z[4], R[9], depth;
  int32_t type, zdim, rdim, active, id;
  bool handled;
};
struct MeasQueue {
  int H = 0, MH = 0, i = 0;
  double* rx = nullptr;  // [H][B][xs]
  double* rP = nullptr;  // [H][B][ps]
  std::vector<double> t;  // t_[slot]
  double start_t = NAN;
  std::deque<std::pair<double, std::array<double, 6>>> u;  // front = newest; body-frame inputs
  std::deque<QueuedMeas> z;                                 // front = newest
};

struct viekf_batch {
  viekf_config cfg;
  MeasQueue* q = nullptr;
  double* x_main = nullptr;  // the allocations kp.x / kp.P pointed to before the ring took over
  double* P_main = nullptr;
  int64_t B = 0;
  int32_t N = 0;
  int device = 0;
  int sm_count = 0;
  int group = 0;
  cudaStream_t stream = nullptr;
  KParams kp;
  int64_t launches = 0;
  // scratch for the per-call path
  viekf_event* d_ev1 = nullptr;
  double* d_vals = nullptr;   // [B][8]
  int32_t* d_ids = nullptr;   // [KF_CAP]
  uint8_t* d_res = nullptr;   // [B]
  double* d_x0 = nullptr;     // [B][17]
  // grow-only device staging owned by the handle (read-back unpack buffers, per-event results, the tape of run_tape): no
  // cudaMalloc / cudaFree on the per-call paths once they are warm
  void* d_stage = nullptr;
  size_t stage_cap = 0;
  void* d_res_stage = nullptr;
  size_t res_cap = 0;
  viekf_event* d_tape_ev = nullptr;
  size_t tape_ev_cap = 0;  // events
  int32_t* d_tape_ids = nullptr;
  size_t tape_ids_cap = 0;  // ids
  double* d_tape_vals = nullptr;
  size_t tape_vals_cap = 0;  // doubles
};

// grow-only allocation: *buf holds at least `need` bytes afterwards
static cudaError_t grow(void** buf, size_t* cap, size_t need) {
  if (need <= *cap) return cudaSuccess;
  cudaFree(*buf);
  *buf = nullptr;
  *cap = 0;
  const size_t want = need + need / 4;
  cudaError_t e = cudaMalloc(buf, want);
  if (e == cudaSuccess) *cap = want;
  return e;
}

static int fail(int code, const char* what, cudaError_t e = cudaSuccess) {
  char buf[512];
  if (e != cudaSuccess)
    snprintf(buf, sizeof(buf), "%s: %s", what, cudaGetErrorString(e));
  else
    snprintf(buf, sizeof(buf), "%s", what);
  g_err = buf;
  return code;
}
#define CK(call)                                                    \
  do {                                                              \
    cudaError_t e__ = (call);                                       \
    if (e__ != cudaSuccess) return fail(VIEKF_ERR_CUDA, #call, e__); \
  } while (0)

static void quat_R(const double* q, double* m) {
  const double w = q[0], x = q[1], y = q[2], z = q[3];
  const double wx = w * x, wy = w * y, wz = w * z, xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
  m[0] = 1. - 2. * yy - 2. * zz; m[1] = 2. * xy + 2. * wz; m[2] = 2. * xz - 2. * wy;
  m[3] = 2. * xy - 2. * wz; m[4] = 1. - 2. * xx - 2. * zz; m[5] = 2. * yz + 2. * wx;
  m[6] = 2. * xz + 2. * wy; m[7] = 2. * yz - 2. * wx; m[8] = 1. - 2. * xx - 2. * yy;
}

static void fill_params(viekf_batch* h) {
  KParams& p = h->kp;
  const viekf_config& c = h->cfg;
  memcpy(p.x0, c.x0, sizeof(p.x0));
  memcpy(p.P0, c.P0, sizeof(p.P0));
  memcpy(p.Qx, c.Qx, sizeof(p.Qx));
  memcpy(p.lam, c.lambda, sizeof(p.lam));
  memcpy(p.Qu, c.Qu, sizeof(p.Qu));
  memcpy(p.P0f, c.P0_feat, sizeof(p.P0f));
  memcpy(p.Qxf, c.Qx_feat, sizeof(p.Qxf));
  memcpy(p.lamf, c.lambda_feat, sizeof(p.lamf));
  memcpy(p.cc, c.cam_center, sizeof(p.cc));
  memcpy(p.fl, c.focal_len, sizeof(p.fl));
  memcpy(p.qbc, c.q_b_c, sizeof(p.qbc));
  memcpy(p.pbc, c.p_b_c, sizeof(p.pbc));
  memcpy(p.qbu, c.q_b_u, sizeof(p.qbu));
  quat_R(c.q_b_c, p.Rbc);
  const double* t = c.p_b_c;
  const double sp[9] = {0, -t[2], t[1], t[2], 0, -t[0], -t[1], t[0], 0};
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      double s = 0;
      for (int k = 0; k < 3; k++) s += p.Rbc[i * 3 + k] * sp[k * 3 + j];
      p.RbcSp[i * 3 + j] = s;
    }
  p.min_depth = c.min_depth;
  p.depth_reset = 1.0 / (2.0 * c.min_depth);
  p.kf_thresh = c.keyframe_overlap_threshold;
  p.use_drag = c.use_drag_term;
  p.use_partial = c.use_partial_update;
  p.use_kfreset = c.use_keyframe_reset;
}

extern "C" {

static int queue_reset_dead(viekf_batch* h);

const char* viekf_last_error(void) { return g_err.c_str(); }

void viekf_config_default(viekf_config* c) {
  // reference params/ekf.yaml
  memset(c, 0, sizeof(*c));
  const double x0[17] = {0, 0, -2, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0.1};
  memcpy(c->x0, x0, sizeof(x0));
  for (int i = 0; i < 9; i++) c->P0[i] = 1e-3;
  for (int i = 9; i < 12; i++) c->P0[i] = 2e-1;
  for (int i = 12; i < 15; i++) c->P0[i] = 1e-1;
  c->P0[15] = 1e-6;
  for (int i = 0; i < 9; i++) c->lambda[i] = 1.0;
  for (int i = 9; i < 15; i++) c->lambda[i] = 0.1;
  c->lambda[15] = 0.01;
  const double Qu[6] = {1.0, 1.0, 1.0, 4e-4, 4e-4, 4e-4};
  memcpy(c->Qu, Qu, sizeof(Qu));
  c->P0_feat[0] = 0.01; c->P0_feat[1] = 0.01; c->P0_feat[2] = 0.3;
  c->lambda_feat[0] = 1.0; c->lambda_feat[1] = 1.0; c->lambda_feat[2] = 0.4;
  c->cam_center[0] = 315.83184814453125; c->cam_center[1] = 242.1165771484375;
  c->focal_len[0] = 611.1864013671875; c->focal_len[1] = 611.5557861328125;
  const double qbc[4] = {0.9974335273839892, 0.019768487288642146, -0.03564306227555538, 0.05886541830542371};
  memcpy(c->q_b_c, qbc, sizeof(qbc));
  const double pbc[3] = {0.17363129, -0.02205945, 0.05490228};
  memcpy(c->p_b_c, pbc, sizeof(pbc));
  const double qbu[4] = {0.993760669165504, 0.0497294816014604, 0.0997086508721388, 0.00498959122946198};
  memcpy(c->q_b_u, qbu, sizeof(qbu));
  c->min_depth = 1.5;
  c->keyframe_overlap_threshold = 0.8;
  c->use_drag_term = 1;
  c->use_partial_update = 1;
  c->use_keyframe_reset = 1;
}

// VIEKF::load (vi_ekf.cpp:101-155) for the flat YAML of params/ekf.yaml: `key: scalar` or `key: [v, v, ...]` (a list may span
// lines), `#` starts a comment.
static bool yaml_flat(const char* path, std::vector<std::pair<std::string, std::vector<std::string>>>& out) {
  FILE* f = fopen(path, "r");
  if (!f) return false;
  std::string text;
  char buf[4096];
  size_t n;
  while ((n = fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, n);
  fclose(f);
  std::string clean;  // comments removed
  bool in_comment = false;
  for (char ch : text) {
    if (ch == '#') in_comment = true;
    if (ch == '\n') in_comment = false;
    if (!in_comment) clean.push_back(ch);
  }
  size_t pos = 0;
  auto trim = [](std::string v) {
    const size_t a = v.find_first_not_of(" \t\r\n\"'"), b = v.find_last_not_of(" \t\r\n\"'");
    return a == std::string::npos ? std::string() : v.substr(a, b - a + 1);
  };
  while (pos < clean.size()) {
    size_t eol = clean.find('\n', pos);
    if (eol == std::string::npos) eol = clean.size();
    std::string line = clean.substr(pos, eol - pos);
    pos = eol + 1;
    const size_t colon = line.find(':');
    if (colon == std::string::npos) continue;
    const std::string key = trim(line.substr(0, colon));
    std::string val = line.substr(colon + 1);
    if (key.empty()) continue;
    std::vector<std::string> items;
    const size_t lb = val.find('[');
    if (lb != std::string::npos) {
      while (val.find(']') == std::string::npos && pos < clean.size()) {  // the list continues on the next lines
        size_t e2 = clean.find('\n', pos);
        if (e2 == std::string::npos) e2 = clean.size();
        val += " " + clean.substr(pos, e2 - pos);
        pos = e2 + 1;
      }
      const size_t rb = val.find(']');
      std::string body = val.substr(lb + 1, (rb == std::string::npos ? val.size() : rb) - lb - 1);
      size_t p0 = 0;
      while (p0 <= body.size()) {
        size_t c = body.find(',', p0);
        if (c == std::string::npos) c = body.size();
        const std::string it = trim(body.substr(p0, c - p0));
        if (!it.empty()) items.push_back(it);
        p0 = c + 1;
      }
    } else {
      const std::string it = trim(val);
      if (!it.empty()) items.push_back(it);
    }
    out.push_back({key, items});
  }
  return true;
}

int viekf_config_load_yaml(const char* path, viekf_config* c) {
  if (!path || !c) return fail(VIEKF_ERR_ARG, "viekf_config_load_yaml: null argument");
  std::vector<std::pair<std::string, std::vector<std::string>>> kv;
  if (!yaml_flat(path, kv)) return fail(VIEKF_ERR_ARG, "viekf_config_load_yaml: cannot read the file");
  viekf_config_default(c);
  struct Vec {
    const char* key;
    double* dst;
    int n;
  };
  const Vec vecs[] = {{"x0", c->x0, 17}, {"P0", c->P0, 16}, {"P0_feat", c->P0_feat, 3}, {"Qx", c->Qx, 16}, {"Qx_feat", c->Qx_feat, 3},
                      {"Qu", c->Qu, 6}, {"lambda", c->lambda, 16}, {"lambda_feat", c->lambda_feat, 3}, {"cam_center", c->cam_center, 2},
                      {"focal_len", c->focal_len, 2}, {"q_b_c", c->q_b_c, 4}, {"p_b_c", c->p_b_c, 3}, {"q_b_u", c->q_b_u, 4},
                      {"min_depth", &c->min_depth, 1}, {"keyframe_overlap_threshold", &c->keyframe_overlap_threshold, 1}};
  auto as_bool = [](const std::string& v) { return v == "true" || v == "True" || v == "1" || v == "yes"; };
  for (auto& e : kv) {
    for (const Vec& v : vecs)
      if (e.first == v.key) {
        if ((int)e.second.size() != v.n) {
          g_err = "viekf_config_load_yaml: key '" + e.first + "' has " + std::to_string(e.second.size()) + " values, expected " + std::to_string(v.n);
          return VIEKF_ERR_ARG;
        }
        for (int i = 0; i < v.n; i++) v.dst[i] = atof(e.second[i].c_str());
      }
    if (e.second.size() == 1) {
      if (e.first == "use_drag_term") c->use_drag_term = as_bool(e.second[0]);
      if (e.first == "use_partial_update") c->use_partial_update = as_bool(e.second[0]);
      if (e.first == "use_keyframe_reset") c->use_keyframe_reset = as_bool(e.second[0]);
    }
  }
  return VIEKF_OK;
}

int viekf_batch_create(const viekf_config* cfg, int64_t B, int32_t N, int32_t device, viekf_batch** out) {
  if (!cfg || !out || B <= 0 || N < 0 || N > 64) return fail(VIEKF_ERR_ARG, "viekf_batch_create: bad argument (need B>0, 0<=N<=64)");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= 0) return fail(VIEKF_ERR_NODEVICE, "viekf_batch_create: no CUDA device (there is no CPU fallback)", e);
  if (device < 0 || device >= ndev) return fail(VIEKF_ERR_ARG, "viekf_batch_create: bad device ordinal");
  CK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  viekf_batch* h = new viekf_batch();
  h->cfg = *cfg;
  h->B = B;
  h->N = N;
  h->device = device;
  h->sm_count = prop.multiProcessorCount;
  h->group = vk::group_size_for(N);
  if (const char* g = getenv("VIEKF_GROUP")) {
    const int v = atoi(g);
    if (v == 32 || v == 64 || v == 128 || v == 192 || v == 256) h->group = v;
    if (v == vk::GROUP_TILE && vk::tile_warps_for(N) > 0) h->group = v;
    if (v == vk::GROUP_OCTET && N == 0) h->group = v;
    if (v == vk::GROUP_FRAG && N >= 1 && N <= 8) h->group = v;
  }
  const size_t need = (h->group == vk::GROUP_OCTET) ? vk::octet_smem_bytes() : (h->group == vk::GROUP_FRAG) ? vk::frag_smem_bytes(N) : (h->group == vk::GROUP_TILE) ? vk::tile_smem_bytes(N)
                                                   : vk::smem_bytes_per_filter(N) * vk::smem_filters_per_cta(h->group);
  if (need > (size_t)prop.sharedMemPerBlockOptin) {
    delete h;
    return fail(VIEKF_ERR_UNSUPPORTED, "viekf_batch_create: N too large for the shared-memory resident covariance");
  }
  memset(&h->kp, 0, sizeof(h->kp));
  fill_params(h);
  KParams& p = h->kp;
  p.N = N;
  p.B = B;
  p.inst_off = 0;
  const int nmax = 16 + 3 * N;
  p.xs = (17 + 5 * N + 1) & ~1;
  p.ps = (nmax * (nmax + 1) / 2 + 1) & ~1;
  auto cleanup = [&](int code, const char* what, cudaError_t ce) {
    viekf_batch_destroy(h);
    return fail(code, what, ce);
  };
  if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) return cleanup(VIEKF_ERR_CUDA, "cudaStreamCreate", e);
  if ((e = cudaMalloc(&p.x, sizeof(double) * p.xs * B)) != cudaSuccess) return cleanup(VIEKF_ERR_NOMEM, "cudaMalloc x", e);
  if ((e = cudaMalloc(&p.P, sizeof(double) * p.ps * B)) != cudaSuccess) return cleanup(VIEKF_ERR_NOMEM, "cudaMalloc P", e);
  if ((e = cudaMalloc(&p.meta, sizeof(FilterMeta) * B)) != cudaSuccess) return cleanup(VIEKF_ERR_NOMEM, "cudaMalloc meta", e);
  if ((e = cudaMalloc(&p.ids, sizeof(int32_t) * (N > 0 ? N : 1) * B)) != cudaSuccess) return cleanup(VIEKF_ERR_NOMEM, "cudaMalloc ids", e);
  if ((e = cudaMalloc(&p.kf_ids, sizeof(int32_t) * vk::KF_CAP * B)) != cudaSuccess) return cleanup(VIEKF_ERR_NOMEM, "cudaMalloc kf_ids", e);
  if ((e = cudaMalloc(&h->d_ev1, sizeof(viekf_event))) != cudaSuccess) return cleanup(VIEKF_ERR_NOMEM, "cudaMalloc ev", e);
  if ((e = cudaMalloc(&h->d_ids, sizeof(int32_t) * vk::KF_CAP)) != cudaSuccess) return cleanup(VIEKF_ERR_NOMEM, "cudaMalloc ids scratch", e);
  if ((e = cudaMemsetAsync(p.x, 0, sizeof(double) * p.xs * B, h->stream)) != cudaSuccess) return cleanup(VIEKF_ERR_CUDA, "memset x", e);
  if ((e = cudaMemsetAsync(p.ids, 0xff, sizeof(int32_t) * (N > 0 ? N : 1) * B, h->stream)) != cudaSuccess) return cleanup(VIEKF_ERR_CUDA, "memset ids", e);
  if ((e = vk::launch_reset(p, h->stream)) != cudaSuccess) return cleanup(VIEKF_ERR_CUDA, "reset kernel", e);
  if ((e = cudaStreamSynchronize(h->stream)) != cudaSuccess) return cleanup(VIEKF_ERR_CUDA, "reset sync", e);
  h->launches = 0;
  *out = h;
  return VIEKF_OK;
}

void viekf_batch_destroy(viekf_batch* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  if (h->q) {
    cudaFree(h->q->rx);
    cudaFree(h->q->rP);
    h->kp.x = h->x_main;
    h->kp.P = h->P_main;
    delete h->q;
  }
  cudaFree(h->kp.x);
  cudaFree(h->kp.P);
  cudaFree(h->kp.meta);
  cudaFree(h->kp.ids);
  cudaFree(h->kp.kf_ids);
  cudaFree(h->kp.stats);
  cudaFree(h->d_ev1);
  cudaFree(h->d_vals);
  cudaFree(h->d_ids);
  cudaFree(h->d_res);
  cudaFree(h->d_x0);
  cudaFree(h->d_stage);
  cudaFree(h->d_res_stage);
  cudaFree(h->d_tape_ev);
  cudaFree(h->d_tape_ids);
  cudaFree(h->d_tape_vals);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

int viekf_batch_reset(viekf_batch* h) {
  if (!h) return fail(VIEKF_ERR_ARG, "null handle");
  CK(cudaSetDevice(h->device));
  if (MeasQueue* q = h->q) {  // back to slot 0 of a zeroed ring with empty deques
    CK(cudaMemsetAsync(q->rx, 0, sizeof(double) * h->kp.xs * h->B * q->H, h->stream));
    CK(cudaMemsetAsync(q->rP, 0, sizeof(double) * h->kp.ps * h->B * q->H, h->stream));
    q->i = 0;
    h->kp.x = q->rx;
    h->kp.P = q->rP;
    q->t.assign(q->H, NAN);
    q->start_t = NAN;
    q->u.clear();
    q->z.clear();
  }
  CK(vk::launch_reset(h->kp, h->stream));
  if (h->q)
    if (int rc = queue_reset_dead(h)) return rc;
  CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

int viekf_batch_set_instance_offset(viekf_batch* h, int64_t first) {
  if (!h || first < 0) return fail(VIEKF_ERR_ARG, "bad argument");
  h->kp.inst_off = first;
  return VIEKF_OK;
}

int viekf_batch_set_x0(viekf_batch* h, const double* x0, int32_t broadcast) {
  if (!h || !x0) return fail(VIEKF_ERR_ARG, "bad argument");
  CK(cudaSetDevice(h->device));
  if (!h->d_x0) CK(cudaMalloc(&h->d_x0, sizeof(double) * 17 * h->B));
  CK(cudaMemcpyAsync(h->d_x0, x0, sizeof(double) * 17 * (broadcast ? 1 : h->B), cudaMemcpyHostToDevice, h->stream));
  CK(vk::launch_set_x0(h->kp, h->d_x0, broadcast, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

int viekf_batch_perturb_x0(viekf_batch* h, uint64_t seed) {
  if (!h) return fail(VIEKF_ERR_ARG, "null handle");
  CK(cudaSetDevice(h->device));
  CK(vk::launch_perturb(h->kp, seed, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

static int check_range(viekf_batch* h, int64_t first, int64_t count, const void* ptr) {
  if (!h || !ptr || first < 0 || count < 0 || first + count > h->B) return fail(VIEKF_ERR_ARG, "bad instance range or null pointer");
  return VIEKF_OK;
}

int viekf_batch_get_state(viekf_batch* h, int64_t first, int64_t count, double* x) {
  if (int rc = check_range(h, first, count, x)) return rc;
  if (count == 0) return VIEKF_OK;
  CK(cudaSetDevice(h->device));
  const size_t bytes = sizeof(double) * (17 + 5 * h->N) * count;
  if (grow(&h->d_stage, &h->stage_cap, bytes) != cudaSuccess) return fail(VIEKF_ERR_NOMEM, "get_state: staging buffer");
  double* d = (double*)h->d_stage;
  cudaError_t e = vk::launch_unpack_state(h->kp, first, count, d, h->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(x, d, bytes, cudaMemcpyDeviceToHost, h->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) return fail(VIEKF_ERR_CUDA, "get_state", e);
  return VIEKF_OK;
}

int viekf_batch_get_covariance(viekf_batch* h, int64_t first, int64_t count, double* P) {
  if (int rc = check_range(h, first, count, P)) return rc;
  if (count == 0) return VIEKF_OK;
  CK(cudaSetDevice(h->device));
  const int64_t n = 16 + 3 * h->N;
  // unpack in slabs so a large batch never needs a full dense copy on the device
  const int64_t slab = std::max<int64_t>(1, (int64_t)(256ll << 20) / (int64_t)(sizeof(double) * n * n));
  if (grow(&h->d_stage, &h->stage_cap, sizeof(double) * n * n * std::min(slab, count)) != cudaSuccess)
    return fail(VIEKF_ERR_NOMEM, "get_covariance: staging buffer");
  double* d = (double*)h->d_stage;
  cudaError_t e = cudaSuccess;
  for (int64_t o = 0; o < count && e == cudaSuccess; o += slab) {
    const int64_t c = std::min(slab, count - o);
    e = vk::launch_unpack_cov(h->kp, first + o, c, d, h->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(P + o * n * n, d, sizeof(double) * n * n * c, cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  }
  if (e != cudaSuccess) return fail(VIEKF_ERR_CUDA, "get_covariance", e);
  return VIEKF_OK;
}

int viekf_batch_get_covariance_diagonal(viekf_batch* h, int64_t first, int64_t count, double* dg) {
  if (int rc = check_range(h, first, count, dg)) return rc;
  if (count == 0) return VIEKF_OK;
  CK(cudaSetDevice(h->device));
  const size_t bytes = sizeof(double) * (16 + 3 * h->N) * count;
  if (grow(&h->d_stage, &h->stage_cap, bytes) != cudaSuccess) return fail(VIEKF_ERR_NOMEM, "get_covariance_diagonal: staging buffer");
  CK(vk::launch_unpack_diag(h->kp, first, count, (double*)h->d_stage, h->stream));
  CK(cudaMemcpyAsync(dg, h->d_stage, bytes, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

// ---- feature read-back (vi_ekf.cpp:210-260): from the unpacked state and the id lists, on the host ----
namespace {
struct FeatView {
  std::vector<double> x;
  std::vector<int32_t> len, ids;
  int mx = 0;
};
void host_zeta(const double* q, double* z) {  // q.rota(e_z) = R^T e_z: the third row of the passive R (quat.h:306-322)
  const double w = q[0], x = q[1], y = q[2], zz = q[3];
  z[0] = 2.0 * x * zz + 2.0 * w * y;
  z[1] = 2.0 * y * zz - 2.0 * w * x;
  z[2] = 1.0 - 2.0 * x * x - 2.0 * y * y;
}
}  // namespace
static int feat_view(viekf_batch* h, int64_t first, int64_t count, FeatView& v) {
  v.mx = 17 + 5 * h->N;
  v.x.resize((size_t)v.mx * count);
  v.len.resize((size_t)count);
  v.ids.resize((size_t)std::max(h->N, 1) * count);
  if (int rc = viekf_batch_get_state(h, first, count, v.x.data())) return rc;
  return viekf_batch_get_features(h, first, count, v.len.data(), v.ids.data());
}

int viekf_batch_get_depths(viekf_batch* h, int64_t first, int64_t count, double* depths) {
  if (int rc = check_range(h, first, count, depths)) return rc;
  if (count == 0 || h->N == 0) return VIEKF_OK;
  FeatView v;
  if (int rc = feat_view(h, first, count, v)) return rc;
  for (int64_t b = 0; b < count; b++)
    for (int i = 0; i < h->N; i++) depths[b * h->N + i] = (i < v.len[b]) ? 1.0 / v.x[b * v.mx + 17 + 5 * i + 4] : NAN;
  return VIEKF_OK;
}

int viekf_batch_get_qzetas(viekf_batch* h, int64_t first, int64_t count, double* qz) {
  if (int rc = check_range(h, first, count, qz)) return rc;
  if (count == 0 || h->N == 0) return VIEKF_OK;
  FeatView v;
  if (int rc = feat_view(h, first, count, v)) return rc;
  for (int64_t b = 0; b < count; b++)
    for (int i = 0; i < h->N; i++)
      for (int k = 0; k < 4; k++) qz[(b * h->N + i) * 4 + k] = (i < v.len[b]) ? v.x[b * v.mx + 17 + 5 * i + k] : NAN;
  return VIEKF_OK;
}

int viekf_batch_get_zetas(viekf_batch* h, int64_t first, int64_t count, double* zetas) {
  if (int rc = check_range(h, first, count, zetas)) return rc;
  if (count == 0 || h->N == 0) return VIEKF_OK;
  FeatView v;
  if (int rc = feat_view(h, first, count, v)) return rc;
  for (int64_t b = 0; b < count; b++)
    for (int i = 0; i < h->N; i++) {
      double* z = zetas + (b * h->N + i) * 3;
      if (i < v.len[b])
        host_zeta(&v.x[b * v.mx + 17 + 5 * i], z);
      else
        z[0] = z[1] = z[2] = NAN;
    }
  return VIEKF_OK;
}

int viekf_batch_get_depth(viekf_batch* h, int32_t id, int64_t first, int64_t count, double* depth) {
  if (int rc = check_range(h, first, count, depth)) return rc;
  if (count == 0) return VIEKF_OK;
  FeatView v;
  if (int rc = feat_view(h, first, count, v)) return rc;
  for (int64_t b = 0; b < count; b++) {
    depth[b] = NAN;
    for (int i = 0; i < v.len[b]; i++)
      if (v.ids[b * h->N + i] == id) depth[b] = 1.0 / v.x[b * v.mx + 17 + 5 * i + 4];
  }
  return VIEKF_OK;
}

int viekf_batch_get_feat(viekf_batch* h, int32_t id, int64_t first, int64_t count, double* pixel) {
  if (int rc = check_range(h, first, count, pixel)) return rc;
  if (count == 0) return VIEKF_OK;
  FeatView v;
  if (int rc = feat_view(h, first, count, v)) return rc;
  for (int64_t b = 0; b < count; b++) {
    pixel[2 * b] = pixel[2 * b + 1] = NAN;
    for (int i = 0; i < v.len[b]; i++)
      if (v.ids[b * h->N + i] == id) {  // cam_F zeta / (e_z^T zeta) + cam_center (vi_ekf.cpp:254-259)
        double z[3];
        host_zeta(&v.x[b * v.mx + 17 + 5 * i], z);
        pixel[2 * b] = h->cfg.focal_len[0] * z[0] / z[2] + h->cfg.cam_center[0];
        pixel[2 * b + 1] = h->cfg.focal_len[1] * z[1] / z[2] + h->cfg.cam_center[1];
      }
  }
  return VIEKF_OK;
}

int viekf_batch_set_imu_bias(viekf_batch* h, const double* b_g, const double* b_a) {
  if (!h || !b_g || !b_a) return fail(VIEKF_ERR_ARG, "bad argument");
  CK(cudaSetDevice(h->device));
  if (grow(&h->d_stage, &h->stage_cap, 6 * sizeof(double)) != cudaSuccess) return fail(VIEKF_ERR_NOMEM, "set_imu_bias: staging buffer");
  double v[6] = {b_g[0], b_g[1], b_g[2], b_a[0], b_a[1], b_a[2]};
  CK(cudaMemcpyAsync(h->d_stage, v, sizeof(v), cudaMemcpyHostToDevice, h->stream));
  CK(vk::launch_set_bias(h->kp, (const double*)h->d_stage, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

int viekf_batch_set_drag_term(viekf_batch* h, int32_t use_drag_term) {
  if (!h) return fail(VIEKF_ERR_ARG, "null handle");
  h->cfg.use_drag_term = use_drag_term ? 1 : 0;
  h->kp.use_drag = h->cfg.use_drag_term;
  return VIEKF_OK;
}
int viekf_batch_get_drag_term(const viekf_batch* h) { return h ? h->cfg.use_drag_term : 0; }


static int fetch_meta(viekf_batch* h, int64_t first, int64_t count, std::vector<FilterMeta>& m) {
  m.resize((size_t)count);
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->stream));
  CK(cudaMemcpy(m.data(), h->kp.meta + first, sizeof(FilterMeta) * count, cudaMemcpyDeviceToHost));
  return VIEKF_OK;
}

int viekf_batch_get_features(viekf_batch* h, int64_t first, int64_t count, int32_t* len, int32_t* ids) {
  if (int rc = check_range(h, first, count, len)) return rc;
  if (count == 0) return VIEKF_OK;
  std::vector<FilterMeta> m;
  if (int rc = fetch_meta(h, first, count, m)) return rc;
  for (int64_t b = 0; b < count; b++) len[b] = m[b].len;
  if (ids && h->N > 0) {
    CK(cudaMemcpy(ids, h->kp.ids + first * h->N, sizeof(int32_t) * h->N * count, cudaMemcpyDeviceToHost));
    for (int64_t b = 0; b < count; b++)
      for (int i = len[b]; i < h->N; i++) ids[b * h->N + i] = -1;
  }
  return VIEKF_OK;
}

int viekf_batch_get_flags(viekf_batch* h, int64_t first, int64_t count, uint32_t* flags, uint32_t* counters) {
  if (int rc = check_range(h, first, count, flags)) return rc;
  if (count == 0) return VIEKF_OK;
  std::vector<FilterMeta> m;
  if (int rc = fetch_meta(h, first, count, m)) return rc;
  for (int64_t b = 0; b < count; b++) {
    flags[b] = m[b].flags;
    if (counters) {
      counters[3 * b] = m[b].n_gated;
      counters[3 * b + 1] = m[b].n_fix;
      counters[3 * b + 2] = m[b].n_nan_gain;
    }
  }
  return VIEKF_OK;
}

int viekf_batch_get_time(viekf_batch* h, int64_t first, int64_t count, double* t) {
  if (int rc = check_range(h, first, count, t)) return rc;
  if (count == 0) return VIEKF_OK;
  std::vector<FilterMeta> m;
  if (int rc = fetch_meta(h, first, count, m)) return rc;
  for (int64_t b = 0; b < count; b++) t[b] = m[b].t;
  return VIEKF_OK;
}

int viekf_batch_get_global_pose(viekf_batch* h, int64_t first, int64_t count, double* node_t3q4, double* gcov36) {
  if (int rc = check_range(h, first, count, node_t3q4)) return rc;
  if (count == 0) return VIEKF_OK;
  std::vector<FilterMeta> m;
  if (int rc = fetch_meta(h, first, count, m)) return rc;
  for (int64_t b = 0; b < count; b++) {
    for (int i = 0; i < 3; i++) node_t3q4[7 * b + i] = m[b].node_t[i];
    for (int i = 0; i < 4; i++) node_t3q4[7 * b + 3 + i] = m[b].node_q[i];
    if (gcov36) memcpy(gcov36 + 36 * b, m[b].gcov, sizeof(double) * 36);
  }
  return VIEKF_OK;
}

// ---- launching --------------------------------------------------------------------------------------------------
static int run_range(viekf_batch* h, const viekf_dev_tape* t, int64_t ev_first, int64_t ev_count, uint8_t* d_results, int64_t res_first,
                     int64_t res_count) {
  KParams p = h->kp;
  p.ev = t->ev;
  p.n_events = t->n_events;
  p.ev_first = ev_first;
  p.ev_count = ev_count;
  p.id_pool = t->id_pool;
  p.values = t->values;
  p.vstride = t->vstride;
  p.seed = t->seed;
  memcpy(p.bias_range, t->bias_range, sizeof(p.bias_range));
  p.results = d_results;
  p.res_first = res_first;
  p.res_count = res_count;
  CK(vk::launch_tape(p, h->group, h->sm_count, h->stream));
  h->launches++;
  return VIEKF_OK;
}

static int run_single_event(viekf_batch* h, const viekf_event& ev, const double* host_vals, int vdim, int broadcast, uint8_t* host_res) {
  CK(cudaSetDevice(h->device));
  viekf_event e = ev;
  viekf_dev_tape t;
  t.ev = h->d_ev1;
  t.n_events = 1;
  t.id_pool = h->d_ids;
  if (host_vals && !broadcast) {
    if (!h->d_vals) CK(cudaMalloc(&h->d_vals, sizeof(double) * 8 * h->B));
    CK(cudaMemcpyAsync(h->d_vals, host_vals, sizeof(double) * vdim * h->B, cudaMemcpyHostToDevice, h->stream));
    t.values = h->d_vals;
    t.vstride = vdim;
    e.noise = 0;
  } else {
    if (host_vals)
      for (int k = 0; k < vdim && k < 8; k++) e.z[k] = host_vals[k];
    e.noise = -1;
  }
  CK(cudaMemcpyAsync(h->d_ev1, &e, sizeof(e), cudaMemcpyHostToDevice, h->stream));
  uint8_t* dres = nullptr;
  if (host_res) {
    if (!h->d_res) CK(cudaMalloc(&h->d_res, h->B));
    dres = h->d_res;
  }
  if (int rc = run_range(h, &t, 0, 1, dres, 0, dres ? h->B : 0)) return rc;
  if (host_res) CK(cudaMemcpyAsync(host_res, dres, h->B, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

static viekf_event blank_event(int kind) {
  viekf_event e;
  memset(&e, 0, sizeof(e));
  e.kind = kind;
  e.id = -1;
  e.noise = -1;
  return e;
}

// clear_feature / keep_only_features; in queue mode a removal also zeroes the dead region of the current slot, as
// vi_ekf_feat.cpp:68-71 does
static int queue_zero_dead(viekf_batch* h, int len);
static int queue_len(viekf_batch* h, int* len);
static int run_removal_event(viekf_batch* h, const viekf_event& e) {
  int before = 0, after = 0;
  if (h->q)
    if (int rc = queue_len(h, &before)) return rc;
  if (int rc = run_single_event(h, e, nullptr, 0, 1, nullptr)) return rc;
  if (h->q) {
    if (int rc = queue_len(h, &after)) return rc;
    if (after < before) {
      if (int rc = queue_zero_dead(h, after)) return rc;
      CK(cudaStreamSynchronize(h->stream));
    }
  }
  return VIEKF_OK;
}

int viekf_batch_propagate(viekf_batch* h, const double* u, int32_t broadcast, double t) {
  if (!h || !u) return fail(VIEKF_ERR_ARG, "bad argument");
  viekf_event e = blank_event(VIEKF_EV_PROP);
  e.t = t;
  return run_single_event(h, e, u, 6, broadcast, nullptr);
}

int viekf_batch_update(viekf_batch* h, int32_t type, const double* z, int32_t zdim, int32_t broadcast, const double* R, int32_t rdim,
                       int32_t active, int32_t id, double depth, uint8_t* result) {
  if (!h || !z || !R || zdim < 1 || zdim > 4 || rdim < 1 || rdim > 3 || type < 0 || type >= VIEKF_TOTAL_MEAS)
    return fail(VIEKF_ERR_ARG, "viekf_batch_update: bad argument");
  viekf_event e = blank_event(VIEKF_EV_UPD);
  e.type = type;
  e.zdim = zdim;
  e.rdim = rdim;
  e.active = active;
  e.id = id;
  e.z[4] = depth;
  for (int i = 0; i < rdim * rdim; i++) e.R[i] = R[i];
  return run_single_event(h, e, z, zdim, broadcast, result);
}

int viekf_batch_init_feature(viekf_batch* h, const double* pixel, int32_t broadcast, int32_t id, double depth, uint8_t* ok) {
  if (!h || !pixel) return fail(VIEKF_ERR_ARG, "bad argument");
  viekf_event e = blank_event(VIEKF_EV_INIT);
  e.id = id;
  e.z[4] = depth;
  return run_single_event(h, e, pixel, 2, broadcast, ok);
}

int viekf_batch_clear_feature(viekf_batch* h, int32_t id) {
  if (!h) return fail(VIEKF_ERR_ARG, "null handle");
  viekf_event e = blank_event(VIEKF_EV_CLEAR);
  e.id = id;
  return run_removal_event(h, e);
}

int viekf_batch_keep_only_features(viekf_batch* h, const int32_t* ids, int32_t n) {
  if (!h || n < 0 || (n > 0 && !ids)) return fail(VIEKF_ERR_ARG, "bad argument");
  if (n > vk::KF_CAP) return fail(VIEKF_ERR_UNSUPPORTED, "keep_only_features: more than 64 ids");
  CK(cudaSetDevice(h->device));
  if (n > 0) CK(cudaMemcpyAsync(h->d_ids, ids, sizeof(int32_t) * n, cudaMemcpyHostToDevice, h->stream));
  viekf_event e = blank_event(VIEKF_EV_KEEP);
  e.n_ids = n;
  e.ids_off = 0;
  return run_removal_event(h, e);
}

int viekf_batch_keyframe_reset(viekf_batch* h) {
  if (!h) return fail(VIEKF_ERR_ARG, "null handle");
  viekf_event e = blank_event(VIEKF_EV_KFRESET);
  return run_single_event(h, e, nullptr, 0, 1, nullptr);
}

// ---- delayed-measurement queue (vi_ekf_meas.cpp:6-194) ------------------------------------------------------------
__global__ void set_time_kernel(FilterMeta* m, int64_t B, double t) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) m[b].t = t;
}

// q.rota(v) on the host, the formula of viekf_math.cuh rota() (reference quat.h:377-391)
static void host_rota(const double* q, const double* v, double* o) {
  const double b[3] = {q[1], q[2], q[3]};
  const double t[3] = {2.0 * (v[1] * b[2] - v[2] * b[1]), 2.0 * (v[2] * b[0] - v[0] * b[2]), 2.0 * (v[0] * b[1] - v[1] * b[0])};
  o[0] = v[0] - q[0] * t[0] + (t[1] * b[2] - t[2] * b[1]);
  o[1] = v[1] - q[0] * t[1] + (t[2] * b[0] - t[0] * b[2]);
  o[2] = v[2] - q[0] * t[2] + (t[0] * b[1] - t[1] * b[0]);
}

// The reference's P is dense over all NUM_FEATURES slots and init() puts P0_feat on every feature block (vi_ekf.cpp:74-80,
// 137-143); each propagate then adds Qx_feat to the dead diagonals too (A_d = I, G_d = 0 there, vi_ekf.cpp:304) and
// clear_feature zeroes the whole dead region (vi_ekf_feat.cpp:68-71).  The kernels never touch the rows behind the live prefix
// of the packed triangle, so in queue mode -- where a rewind can bring a dead block back to life without an init_feature --
// these rows are kept as the reference would have them.
__global__ void queue_dead_diag_kernel(double* P, int64_t ps, const FilterMeta* meta, int64_t B, int N, double a0, double a1, double a2,
                                       int set) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * N) return;
  const int64_t b = idx / N;
  const int f = (int)(idx % N);
  if (f < meta[b].len) return;
  const double a[3] = {a0, a1, a2};
  for (int k = 0; k < 3; k++) {
    const int64_t i = 16 + 3 * f + k;
    double* d = P + b * ps + i * (i + 1) / 2 + i;
    *d = set ? a[k] : *d + a[k];
  }
}

static int queue_dead_diag(viekf_batch* h, const double* a, int set) {
  if (h->N == 0) return VIEKF_OK;
  const int64_t n = h->B * h->N;
  queue_dead_diag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(h->kp.P, h->kp.ps, h->kp.meta, h->B, h->N, a[0], a[1], a[2], set);
  CK(cudaGetLastError());
  return VIEKF_OK;
}

// zero x and packed P behind the live prefix of `len` features in the current slot (vi_ekf_feat.cpp:68-71)
static int queue_zero_dead(viekf_batch* h, int len) {
  const int n = 16 + 3 * len, m = 17 + 5 * len;
  const size_t p0 = (size_t)n * (n + 1) / 2;
  if ((size_t)h->kp.ps > p0)
    CK(cudaMemset2DAsync(h->kp.P + p0, sizeof(double) * h->kp.ps, 0, sizeof(double) * (h->kp.ps - p0), h->B, h->stream));
  if (h->kp.xs > m) CK(cudaMemset2DAsync(h->kp.x + m, sizeof(double) * h->kp.xs, 0, sizeof(double) * (h->kp.xs - m), h->B, h->stream));
  return VIEKF_OK;
}

static int queue_reset_dead(viekf_batch* h) {  // after launch_reset: every feature block dead, P0_feat on its diagonal
  if (int rc = queue_zero_dead(h, 0)) return rc;
  return queue_dead_diag(h, h->cfg.P0_feat, 1);
}

static int queue_len(viekf_batch* h, int* len) {
  std::vector<FilterMeta> m;
  if (int rc = fetch_meta(h, 0, 1, m)) return rc;
  *len = m[0].len;
  return VIEKF_OK;
}

static void queue_select_slot(viekf_batch* h, int slot) {
  MeasQueue* q = h->q;
  q->i = slot;
  h->kp.x = q->rx + (size_t)slot * h->kp.xs * h->B;
  h->kp.P = q->rP + (size_t)slot * h->kp.ps * h->B;
}

int viekf_batch_queue_enable(viekf_batch* h, int32_t len_state_hist, int32_t len_meas_hist) {
  if (!h || len_state_hist < 2 || len_meas_hist < 1) return fail(VIEKF_ERR_ARG, "queue_enable: bad argument");
  if (h->q) return fail(VIEKF_ERR_ARG, "queue_enable: already enabled");
  CK(cudaSetDevice(h->device));
  std::vector<FilterMeta> m;
  if (int rc = fetch_meta(h, 0, 1, m)) return rc;
  if (m[0].started) return fail(VIEKF_ERR_ARG, "queue_enable: must precede the first propagate");
  MeasQueue* q = new MeasQueue();
  q->H = len_state_hist;
  q->MH = len_meas_hist;
  q->t.assign(q->H, NAN);
  const size_t xb = sizeof(double) * h->kp.xs * h->B, pb = sizeof(double) * h->kp.ps * h->B;
  cudaError_t e = cudaMalloc(&q->rx, xb * q->H);
  if (e == cudaSuccess) e = cudaMalloc(&q->rP, pb * q->H);
  if (e != cudaSuccess) {
    cudaFree(q->rx);
    cudaFree(q->rP);
    delete q;
    cudaGetLastError();
    return fail(VIEKF_ERR_NOMEM, "queue_enable: the history ring does not fit in device memory", e);
  }
  // the reference's slots start as zeros except slot 0 (vi_ekf.cpp:18-155)
  CK(cudaMemsetAsync(q->rx, 0, xb * q->H, h->stream));
  CK(cudaMemsetAsync(q->rP, 0, pb * q->H, h->stream));
  CK(cudaMemcpyAsync(q->rx, h->kp.x, xb, cudaMemcpyDeviceToDevice, h->stream));
  CK(cudaMemcpyAsync(q->rP, h->kp.P, pb, cudaMemcpyDeviceToDevice, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  h->x_main = h->kp.x;
  h->P_main = h->kp.P;
  h->q = q;
  queue_select_slot(h, 0);
  if (int rc = queue_zero_dead(h, m[0].len)) return rc;
  if (int rc = queue_dead_diag(h, h->cfg.P0_feat, 1)) return rc;
  CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

// propagate_state(u, t, save_input) on the ring (vi_ekf.cpp:262-318)
static int queue_propagate_impl(viekf_batch* h, const double* u, double t, bool save_input) {
  MeasQueue* q = h->q;
  if (save_input) {
    std::array<double, 6> ub;
    host_rota(h->cfg.q_b_u, u, ub.data());
    host_rota(h->cfg.q_b_u, u + 3, ub.data() + 3);
    q->u.push_front({t, ub});
  }
  viekf_event e = blank_event(VIEKF_EV_PROP);
  e.t = t;
  bool advance = false;
  if (std::isnan(q->start_t)) {  // first call latches the time (vi_ekf.cpp:274-279)
    q->start_t = t;
    q->t[q->i] = t;
  } else {
    const double dt = t - q->t[q->i];
    advance = !(fabs(dt) < 1e-6) && !(dt < 0);  // vi_ekf.cpp:281-290
  }
  if (advance) {
    // x_[ip] receives the live entries only, P_[ip] is assigned whole (vi_ekf.cpp:301,304): copy the live prefix of x and all
    // of P from slot i to slot ip and propagate in place there; what slot ip held in x beyond the prefix stays, as in the
    // reference.
    std::vector<FilterMeta> m;
    if (int rc = fetch_meta(h, 0, 1, m)) return rc;
    const int len = m[0].len;
    const int ip = (q->i + 1) % q->H;
    const double* sx = h->kp.x;
    const double* sP = h->kp.P;
    queue_select_slot(h, ip);
    CK(cudaMemcpy2DAsync(h->kp.x, sizeof(double) * h->kp.xs, sx, sizeof(double) * h->kp.xs, sizeof(double) * (17 + 5 * len), h->B,
                         cudaMemcpyDeviceToDevice, h->stream));
    CK(cudaMemcpyAsync(h->kp.P, sP, sizeof(double) * h->kp.ps * h->B, cudaMemcpyDeviceToDevice, h->stream));  // dead rows included
    const double* qf = h->cfg.Qx_feat;
    if (qf[0] != 0.0 || qf[1] != 0.0 || qf[2] != 0.0)
      if (int rc = queue_dead_diag(h, qf, 0)) return rc;
    q->t[ip] = t;
  }
  return run_single_event(h, e, u, 6, 1, nullptr);
}

int viekf_batch_queue_propagate(viekf_batch* h, const double* u, double t, int32_t save_input) {
  if (!h || !u) return fail(VIEKF_ERR_ARG, "bad argument");
  if (!h->q) return fail(VIEKF_ERR_ARG, "queue_propagate: call viekf_batch_queue_enable first");
  CK(cudaSetDevice(h->device));
  return queue_propagate_impl(h, u, t, save_input != 0);
}

int viekf_batch_add_measurement(viekf_batch* h, double t, int32_t type, const double* z, int32_t zdim, const double* R, int32_t rdim,
                                int32_t active, int32_t id, double depth, int32_t* result) {
  if (!h || !z || !R || !result || zdim < 1 || zdim > 4 || rdim < 1 || rdim > 3 || type < 0 || type >= VIEKF_TOTAL_MEAS)
    return fail(VIEKF_ERR_ARG, "viekf_batch_add_measurement: bad argument");
  MeasQueue* q = h->q;
  if (!q) return fail(VIEKF_ERR_ARG, "add_measurement: call viekf_batch_queue_enable first");
  CK(cudaSetDevice(h->device));
  if (t < q->start_t) {  // vi_ekf_meas.cpp:133-134
    *result = VIEKF_MEAS_INVALID;
    return VIEKF_OK;
  }
  for (int i = 0; i < zdim; i++)
    if (z[i] != z[i]) {  // :136-137
      *result = VIEKF_MEAS_NAN;
      return VIEKF_OK;
    }
  if (type == VIEKF_FEAT && id >= 0) {  // :140-147: a new id is initialised now, in the current slot, and not queued
    std::vector<FilterMeta> m;
    if (int rc = fetch_meta(h, 0, 1, m)) return rc;
    std::vector<int32_t> ids(h->N > 0 ? h->N : 1, -1);
    if (h->N > 0) CK(cudaMemcpy(ids.data(), h->kp.ids, sizeof(int32_t) * h->N, cudaMemcpyDeviceToHost));
    bool tracked = false;
    for (int i = 0; i < m[0].len; i++) tracked |= (ids[i] == id);
    if (!tracked) {
      viekf_event e = blank_event(VIEKF_EV_INIT);
      e.id = id;
      e.z[4] = depth;
      if (int rc = run_single_event(h, e, z, 2, 1, nullptr)) return rc;
      *result = VIEKF_MEAS_NEW_FEATURE;
      return VIEKF_OK;
    }
  }
  auto it = q->z.begin();
  while (it != q->z.end() && !(it->t < t)) ++it;  // :150-156: before the first strictly older entry
  QueuedMeas qm;
  memset(&qm, 0, sizeof(qm));
  qm.t = t;
  qm.type = type;
  qm.zdim = zdim;
  qm.rdim = rdim;
  qm.active = active;
  qm.id = id;
  qm.depth = depth;
  qm.handled = false;
  for (int i = 0; i < zdim; i++) qm.z[i] = z[i];
  for (int i = 0; i < rdim * rdim; i++) qm.R[i] = R[i];
  q->z.insert(it, qm);
  *result = VIEKF_MEAS_SUCCESS;
  return VIEKF_OK;
}

int viekf_batch_handle_measurements(viekf_batch* h, int32_t* upd_types, int32_t* upd_ids, uint8_t* upd_results, int32_t cap,
                                    int32_t* n_updates) {
  if (!h || cap < 0) return fail(VIEKF_ERR_ARG, "bad argument");
  MeasQueue* q = h->q;
  if (!q) return fail(VIEKF_ERR_ARG, "handle_measurements: call viekf_batch_queue_enable first");
  CK(cudaSetDevice(h->device));
  int n_upd = 0;
  if (n_updates) *n_updates = 0;
  if (q->z.empty()) return VIEKF_OK;
  // oldest pending measurement (vi_ekf_meas.cpp:13-22)
  auto z_it = q->z.end() - 1;
  while (z_it->handled && z_it != q->z.begin()) --z_it;
  if (z_it == q->z.begin() && z_it->handled) return VIEKF_OK;
  if (q->u.empty() || z_it->t > q->u.begin()->first) return VIEKF_OK;  // :24-28 measurement from the future
  // newest input strictly older than the measurement (:31-39)
  auto u_it = q->u.begin();
  while (u_it != q->u.end() && !(z_it->t > u_it->first)) ++u_it;
  if (u_it == q->u.end()) return VIEKF_OK;  // the reference dereferences end() here; nothing to rewind to
  // rewind the ring to the newest state not after that input (:45-63)
  int k = q->H;
  int slot = -1;
  while (k > 0) {
    const int s = (q->i + k) % q->H;
    if (q->t[s] <= u_it->first) {
      slot = s;
      break;
    }
    k--;
  }
  if (getenv("VIEKF_QUEUE_DEBUG"))
    fprintf(stderr, "[queue] z.t=%.6f type=%d u.t=%.6f i=%d t[i]=%.6f -> slot %d nz=%zu nu=%zu\n", z_it->t, z_it->type, u_it->first, q->i,
            q->t[q->i], slot, q->z.size(), q->u.size());
  if (slot < 0) {  // older than the ring: the measurement is dropped
    q->z.erase(z_it);
    return VIEKF_OK;
  }
  if (slot != q->i) {
    queue_select_slot(h, slot);
    set_time_kernel<<<(unsigned)((h->B + 255) / 256), 256, 0, h->stream>>>(h->kp.meta, h->B, q->t[slot]);
    CK(cudaGetLastError());
  }
  auto apply_update = [&](QueuedMeas& m) -> int {
    m.handled = true;
    viekf_event e = blank_event(VIEKF_EV_UPD);
    e.type = m.type;
    e.zdim = m.zdim;
    e.rdim = m.rdim;
    e.active = m.active;
    e.id = m.id;
    e.z[4] = m.depth;
    e.aux = 2;  // update() without add_measurement's bookkeeping: a feature dropped since it was queued is MEAS_INVALID, not re-initialised
    for (int i = 0; i < m.rdim * m.rdim; i++) e.R[i] = m.R[i];
    uint8_t* out = (upd_results && n_upd < cap) ? upd_results + (size_t)n_upd * h->B : nullptr;
    if (int rc = run_single_event(h, e, m.z, m.zdim, 1, out)) return rc;
    if (n_upd < cap) {
      if (upd_types) upd_types[n_upd] = m.type;
      if (upd_ids) upd_ids[n_upd] = m.id;
    }
    n_upd++;
    return VIEKF_OK;
  };
  // replay (:73-117); the stored inputs are body-frame and propagate_state rotates them again, as the reference does
  --u_it;
  while (u_it != q->u.begin()) {
    while (z_it->t <= u_it->first) {
      if (q->t[q->i] < z_it->t)
        if (int rc = queue_propagate_impl(h, u_it->second.data(), z_it->t, false)) return rc;
      if (!z_it->handled)
        if (int rc = apply_update(*z_it)) return rc;
      if (z_it != q->z.begin()) {
        --z_it;
        while (u_it->first < z_it->t && u_it != q->u.begin()) {
          if (int rc = queue_propagate_impl(h, u_it->second.data(), u_it->first, false)) return rc;
          --u_it;
        }
      } else {
        while (u_it != q->u.begin()) {
          if (int rc = queue_propagate_impl(h, u_it->second.data(), u_it->first, false)) return rc;
          --u_it;
        }
        break;
      }
    }
  }
  if (int rc = queue_propagate_impl(h, u_it->second.data(), u_it->first, false)) return rc;
  while ((int)q->z.size() > q->MH) q->z.pop_back();
  while ((int)q->u.size() > q->H) q->u.pop_back();
  if (n_updates) *n_updates = n_upd;
  return VIEKF_OK;
}

static int validate_tape(const viekf_tape* t) {
  if (!t || t->n_events < 0 || (t->n_events > 0 && !t->events)) return fail(VIEKF_ERR_ARG, "tape: null events");
  for (int64_t i = 0; i < t->n_events; i++) {
    const viekf_event& e = t->events[i];
    if (e.kind < 0 || e.kind > VIEKF_EV_KFRESET) return fail(VIEKF_ERR_ARG, "tape: unknown event kind");
    if (e.kind == VIEKF_EV_UPD && (e.zdim < 1 || e.zdim > 4 || e.rdim < 1 || e.rdim > 3 || e.type < 0 || e.type >= VIEKF_TOTAL_MEAS))
      return fail(VIEKF_ERR_ARG, "tape: bad UPD event");
    if (e.kind == VIEKF_EV_KEEP) {
      if (e.n_ids < 0 || e.ids_off < 0 || e.ids_off + e.n_ids > t->n_ids || (e.n_ids > 0 && !t->id_pool)) return fail(VIEKF_ERR_ARG, "tape: KEEP ids out of range");
      if (e.n_ids > vk::KF_CAP) return fail(VIEKF_ERR_UNSUPPORTED, "tape: KEEP with more than 64 ids");
    }
    if (e.noise >= 0 && t->values && 8 * ((int64_t)e.noise + 1) > t->values_stride) return fail(VIEKF_ERR_ARG, "tape: value block beyond values_stride");
  }
  return VIEKF_OK;
}

int viekf_batch_upload_tape(viekf_batch* h, const viekf_tape* tape, viekf_dev_tape** out) {
  if (!h || !out) return fail(VIEKF_ERR_ARG, "bad argument");
  if (int rc = validate_tape(tape)) return rc;
  CK(cudaSetDevice(h->device));
  viekf_dev_tape* d = new viekf_dev_tape();
  d->n_events = tape->n_events;
  d->seed = tape->seed;
  memcpy(d->bias_range, tape->imu_bias_range, sizeof(d->bias_range));
  cudaError_t e = cudaSuccess;
  if (tape->n_events > 0) {
    e = cudaMalloc(&d->ev, sizeof(viekf_event) * tape->n_events);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d->ev, tape->events, sizeof(viekf_event) * tape->n_events, cudaMemcpyHostToDevice, h->stream);
  }
  if (e == cudaSuccess && tape->n_ids > 0 && tape->id_pool) {
    e = cudaMalloc(&d->id_pool, sizeof(int32_t) * tape->n_ids);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d->id_pool, tape->id_pool, sizeof(int32_t) * tape->n_ids, cudaMemcpyHostToDevice, h->stream);
  }
  if (e == cudaSuccess && tape->values && tape->values_stride > 0) {
    d->vstride = tape->values_stride;
    e = cudaMalloc(&d->values, sizeof(double) * tape->values_stride * h->B);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(d->values, tape->values, sizeof(double) * tape->values_stride * h->B, cudaMemcpyHostToDevice, h->stream);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) {
    viekf_batch_free_tape(h, d);
    return fail(e == cudaErrorMemoryAllocation ? VIEKF_ERR_NOMEM : VIEKF_ERR_CUDA, "upload_tape", e);
  }
  *out = d;
  return VIEKF_OK;
}

void viekf_batch_free_tape(viekf_batch* h, viekf_dev_tape* t) {
  if (!t) return;
  if (h) {
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
  }
  cudaFree(t->ev);
  cudaFree(t->id_pool);
  cudaFree(t->values);
  delete t;
}

int viekf_batch_run_dev_tape(viekf_batch* h, const viekf_dev_tape* t, int64_t ev_first, int64_t ev_count, int32_t async) {
  if (!h || !t) return fail(VIEKF_ERR_ARG, "bad argument");
  if (ev_count < 0) ev_count = t->n_events - ev_first;
  if (ev_first < 0 || ev_first + ev_count > t->n_events) return fail(VIEKF_ERR_ARG, "event range out of the tape");
  if (ev_count == 0) return VIEKF_OK;
  CK(cudaSetDevice(h->device));
  if (int rc = run_range(h, t, ev_first, ev_count, nullptr, 0, 0)) return rc;
  if (!async) CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

int viekf_batch_run_tape(viekf_batch* h, const viekf_tape* tape, uint8_t* results, int64_t result_first, int64_t result_count) {
  if (!h) return fail(VIEKF_ERR_ARG, "null handle");
  if (results && (result_first < 0 || result_count < 0 || result_first + result_count > h->B)) return fail(VIEKF_ERR_ARG, "bad result range");
  if (int rc = validate_tape(tape)) return rc;
  if (tape->n_events == 0) return VIEKF_OK;
  CK(cudaSetDevice(h->device));
  // the tape goes to grow-only device buffers owned by the handle: one cudaMemcpyAsync per array, no allocation once warm
  viekf_dev_tape d;
  d.n_events = tape->n_events;
  d.seed = tape->seed;
  memcpy(d.bias_range, tape->imu_bias_range, sizeof(d.bias_range));
  {
    size_t cap = h->tape_ev_cap * sizeof(viekf_event);
    if (grow((void**)&h->d_tape_ev, &cap, sizeof(viekf_event) * tape->n_events) != cudaSuccess) return fail(VIEKF_ERR_NOMEM, "run_tape: events");
    h->tape_ev_cap = cap / sizeof(viekf_event);
    d.ev = h->d_tape_ev;
    CK(cudaMemcpyAsync(d.ev, tape->events, sizeof(viekf_event) * tape->n_events, cudaMemcpyHostToDevice, h->stream));
  }
  if (tape->n_ids > 0 && tape->id_pool) {
    size_t cap = h->tape_ids_cap * sizeof(int32_t);
    if (grow((void**)&h->d_tape_ids, &cap, sizeof(int32_t) * tape->n_ids) != cudaSuccess) return fail(VIEKF_ERR_NOMEM, "run_tape: id pool");
    h->tape_ids_cap = cap / sizeof(int32_t);
    d.id_pool = h->d_tape_ids;
    CK(cudaMemcpyAsync(d.id_pool, tape->id_pool, sizeof(int32_t) * tape->n_ids, cudaMemcpyHostToDevice, h->stream));
  }
  if (tape->values && tape->values_stride > 0) {
    size_t cap = h->tape_vals_cap * sizeof(double);
    const size_t need = sizeof(double) * tape->values_stride * h->B;
    if (grow((void**)&h->d_tape_vals, &cap, need) != cudaSuccess) return fail(VIEKF_ERR_NOMEM, "run_tape: values");
    h->tape_vals_cap = cap / sizeof(double);
    d.values = h->d_tape_vals;
    d.vstride = tape->values_stride;
    CK(cudaMemcpyAsync(d.values, tape->values, need, cudaMemcpyHostToDevice, h->stream));
  }
  uint8_t* dres = nullptr;
  const size_t res_bytes = (results && result_count > 0) ? (size_t)result_count * tape->n_events : 0;
  if (res_bytes) {
    if (grow(&h->d_res_stage, &h->res_cap, res_bytes) != cudaSuccess) return fail(VIEKF_ERR_NOMEM, "run_tape: results");
    dres = (uint8_t*)h->d_res_stage;
    CK(cudaMemsetAsync(dres, 0xff, res_bytes, h->stream));
  }
  int rc = run_range(h, &d, 0, tape->n_events, dres, result_first, dres ? result_count : 0);
  cudaError_t e = cudaSuccess;
  if (rc == VIEKF_OK && dres) e = cudaMemcpyAsync(results, dres, res_bytes, cudaMemcpyDeviceToHost, h->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  d.ev = nullptr;  // the buffers belong to the handle
  d.id_pool = nullptr;
  d.values = nullptr;
  if (rc != VIEKF_OK) return rc;
  if (e != cudaSuccess) return fail(VIEKF_ERR_CUDA, "run_tape", e);
  return VIEKF_OK;
}

int viekf_batch_sync(viekf_batch* h) {
  if (!h) return fail(VIEKF_ERR_ARG, "null handle");
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

void* viekf_batch_stream(viekf_batch* h) { return h ? (void*)h->stream : nullptr; }
int64_t viekf_batch_launch_count(const viekf_batch* h) { return h ? h->launches : 0; }

int viekf_batch_stats_reset(viekf_batch* h, int32_t n_buckets) {
  if (!h || n_buckets < 0) return fail(VIEKF_ERR_ARG, "bad argument");
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->stream));
  if (n_buckets != h->kp.n_buckets) {
    cudaFree(h->kp.stats);
    h->kp.stats = nullptr;
    h->kp.n_buckets = 0;
    if (n_buckets > 0) {
      CK(cudaMalloc(&h->kp.stats, sizeof(double) * VIEKF_STAT_WIDTH * n_buckets));
      h->kp.n_buckets = n_buckets;
    }
  }
  if (h->kp.stats) CK(cudaMemsetAsync(h->kp.stats, 0, sizeof(double) * VIEKF_STAT_WIDTH * n_buckets, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

int viekf_batch_stats(viekf_batch* h, double* out, int32_t n_buckets) {
  if (!h || !out || n_buckets < 0 || n_buckets > h->kp.n_buckets) return fail(VIEKF_ERR_ARG, "bad argument");
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->stream));
  if (n_buckets > 0) CK(cudaMemcpy(out, h->kp.stats, sizeof(double) * VIEKF_STAT_WIDTH * n_buckets, cudaMemcpyDeviceToHost));
  return VIEKF_OK;
}

void* viekf_batch_stats_dev_ptr(viekf_batch* h) { return h ? (void*)h->kp.stats : nullptr; }

// ---- read-back without any allocation (no counterpart in the reference) ------
// The unpack kernel of get_state needs device scratch between the engine's layout and the copy to the caller's buffer.  Here
// that scratch is a file-scope device array of fixed size instead of the handle's grow-only cudaMalloc staging, so the call can
// be made where cudaMalloc is forbidden.  Every GPU that loads the library holds the 2 MiB until the process exits.
static __device__ __align__(256) double g_noalloc_scratch[VIEKF_NOALLOC_SCRATCH_DOUBLES];
// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is looked up once per
// device, with that device current, and cached here.
static const int NOALLOC_MAX_DEVICES = 64;
static double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};
static std::mutex g_noalloc_mutex;

static int noalloc_scratch_for(int device, double** out) {
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail(VIEKF_ERR_UNSUPPORTED, "get_state_noalloc: device ordinal beyond 63");
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[device]) {
    double* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);  // the caller has made `device` current
    if (e != cudaSuccess || !ptr) return fail(VIEKF_ERR_CUDA, "get_state_noalloc: cudaGetSymbolAddress(g_noalloc_scratch)", e);
    g_noalloc_ptr[device] = ptr;
  }
  *out = g_noalloc_ptr[device];
  return VIEKF_OK;
}

int viekf_batch_get_state_noalloc(viekf_batch* h, int64_t first, int64_t count, double* x) {
  if (int rc = check_range(h, first, count, x)) return rc;
  if (count == 0) return VIEKF_OK;
  const int64_t need = (int64_t)(17 + 5 * h->N) * count;
  if (need > (int64_t)VIEKF_NOALLOC_SCRATCH_DOUBLES) {
    char buf[256];
    snprintf(buf, sizeof(buf),
             "viekf_batch_get_state_noalloc: %lld instances x %d doubles = %lld doubles exceed the fixed scratch of %d doubles; "
             "read fewer instances per call (at most %lld) or use viekf_batch_get_state",
             (long long)count, 17 + 5 * h->N, (long long)need, (int)VIEKF_NOALLOC_SCRATCH_DOUBLES,
             (long long)(VIEKF_NOALLOC_SCRATCH_DOUBLES / (17 + 5 * h->N)));
    return fail(VIEKF_ERR_ARG, buf);
  }
  CK(cudaSetDevice(h->device));
  double* d = nullptr;
  if (int rc = noalloc_scratch_for(h->device, &d)) return rc;
  // the kernel writes every one of the `need` entries (dead entries as 0), so nothing of the previous call's contents survives
  // in the part that is copied out and nothing has to be cleared first
  CK(vk::launch_unpack_state(h->kp, first, count, d, h->stream));
  CK(cudaMemcpyAsync(x, d, sizeof(double) * need, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return VIEKF_OK;
}

int viekf_fp64_peak_probe(int32_t device, double* tflops) {
  if (!tflops) return fail(VIEKF_ERR_ARG, "null pointer");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return fail(VIEKF_ERR_NODEVICE, "no such CUDA device");
  CK(vk::fp64_probe(device, tflops));
  return VIEKF_OK;
}

}  // extern "C"
