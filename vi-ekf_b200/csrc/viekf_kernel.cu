// viekf_kernel.cu -- persistent tape executor of the VI-EKF batch engine (sm_100a, FP64).
//
// One group of G threads owns one filter instance: its state x and packed-symmetric covariance P are loaded from
// HBM into shared memory once, every event of the tape slice is applied on chip, and x, P go back to HBM once.
// Events follow the in-order semantics of include/viekf_batch.h; each handler names the reference code it replaces:
//   PROP    VIEKF::propagate_state + dynamics + boxplus + fix_depth   (vi_ekf.cpp:262-318, vi_ekf_dyn.cpp:14-135)
//   UPD     add_measurement bookkeeping + VIEKF::update + h_*          (vi_ekf_meas.cpp:130-147,196-395)
//   INIT    init_feature                                               (vi_ekf_feat.cpp:6-47)
//   CLEAR / KEEP  clear_feature / keep_only_features                    (vi_ekf_feat.cpp:50-142)
//   KFRESET keyframe_reset                                              (vi_ekf_kfr.cpp:56-157)
//   STAT    RMSE / NEES partial sums (names only in multirotor_sim/mclogger.h:50-52)
//
// Structure exploited: the discrete transition is [[A_bb,0],[A_fb,blkdiag(A_ff,i)]], so
//   P'_bb = A_bb P_bb A_bb^T,  Y_i = A_fb,i P_bb + A_ff,i P_fb,i,  P'_fb,i = Y_i A_bb^T,
//   P'_ff,ij = Y_i A_fb,j^T + (A_fb,i P_bf,j + A_ff,i P_ff,ij) A_ff,j^T            (+ G_d Qu G_d^T + Qx)
// is evaluated in place block by block, and every H has <= 6 non-zero columns, so the Joseph update
//   P' = (I-KH) P (I-KH)^T + K R K^T = P - K(HP) - (HP)^T K^T + K S K^T
// is a symmetric rank update.  Only the live part n = 16 + 3 len of the MAX_DX system is touched: the dead slots
// of the reference's dense product contribute exact zeros.
#include "viekf_dev.cuh"

namespace vk {


size_t smem_bytes_per_filter(int N) { return (size_t)make_layout(N).total * sizeof(double); }

int group_size_for(int N) {
  if (N == 0) return GROUP_OCTET;
  if (N <= 8) return GROUP_FRAG;  // 1 ... 8 features: covariance in tensor-core fragments (VIEKF_GROUP=32 / 64 / 1 select the older mappings)
  if (tile_warps_for(N) > 0) return GROUP_TILE;
  return 128;
}


// ------------------------------------------------------------------------------------------------------------------
// the persistent tape executor
// ------------------------------------------------------------------------------------------------------------------
#ifndef VIEKF_SMALLG_LOCKSTEP
#define VIEKF_SMALLG_LOCKSTEP 1
#endif
#ifndef VIEKF_G64_CTA
#define VIEKF_G64_CTA 256  // G = 64 (N = 2 ... 6): 4 filters per CTA, 2 CTAs per SM at 128 registers (profiles/r1_v8_smallg_ab.txt)
#define VIEKF_G64_MINB 2
#endif
#ifndef VIEKF_SMALLG_CTA
#define VIEKF_SMALLG_CTA 128  // threads per CTA for G = 32 / 64: VIEKF_SMALLG_CTA / G filters per CTA
#endif
#ifndef VIEKF_SMALLG_MINB
#define VIEKF_SMALLG_MINB 3  // G = 32 / 64 (N <= 6): CTAs of 128 threads per SM the register allocation must admit
#endif
// filters per CTA of the shared-memory executor for a group size: the one constant behind the kernel's launch bounds, its
// launcher and the create-time shared-memory check
int smem_filters_per_cta(int group) {
  if (group >= 128) return 1;
  return (group == 64 ? VIEKF_G64_CTA : VIEKF_SMALLG_CTA) / group;
}

template <int G>
__global__ void __launch_bounds__((G >= 128 ? G : (G == 64 ? VIEKF_G64_CTA : VIEKF_SMALLG_CTA)), (G == 128 ? 3 : (G == 192 ? 3 : (G == 256 ? 3 : (G == 64 ? VIEKF_G64_MINB : VIEKF_SMALLG_MINB))))) viekf_tape_kernel(const __grid_constant__ KParams p) {
  extern __shared__ __align__(16) double smem[];
  constexpr int CTA = (G >= 128 ? G : (G == 64 ? VIEKF_G64_CTA : VIEKF_SMALLG_CTA));
  constexpr int GPC = CTA / G;
  const Layout L = make_layout(p.N);
  Ctx<G> c;
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
  const int nmax = 16 + 3 * p.N;
  for (int i = c.tid; i < nmax; i += G) c.lam[i] = (i < 16) ? p.lam[i] : p.lamf[(i - 16) % 3];

  // G < 128: the GPC filters of a CTA run the shared tape in event lock-step (one CTA barrier per event), so that the warps of
  // an SM stay within one event of each other and share the instruction stream (same reasoning as FPC in viekf_tile.cu)
  constexpr bool LOCK = (GPC > 1) && (VIEKF_SMALLG_LOCKSTEP != 0);
  for (int64_t b0 = (int64_t)blockIdx.x * GPC; b0 < p.B; b0 += (int64_t)gridDim.x * GPC) {
    const int64_t b = b0 + c.gi;
    if (b >= p.B) {  // ragged last round: keep the barrier count of the groups that do have a filter
      if (LOCK)
        for (int64_t e = 0; e < p.ev_count; e++) __syncthreads();
      continue;
    }
    c.b = b;
    const int64_t ginst = p.inst_off + b;
    FilterMeta& M = p.meta[b];
    // ---- load the filter: meta, ids, live part of x and of the packed P ----
    c.sync();
    const int len0 = M.len;
    if (c.tid == 0) {
      c.sc[SC_LEN] = len0; c.sc[SC_NEXT] = M.next_id; c.sc[SC_STARTED] = M.started; c.sc[SC_FLAGS] = (int)M.flags;
      c.sc[SC_NGATED] = (int)M.n_gated; c.sc[SC_NFIX] = (int)M.n_fix; c.sc[SC_NNAN] = (int)M.n_nan_gain; c.sc[SC_NKF] = M.n_kf;
      *c.tcur = M.t;
      for (int k = 0; k < 6; k++) c.sm[SM_UB + k] = M.ub[k];
    }
    for (int i = c.tid; i < len0; i += G) c.ids[i] = p.ids[b * (p.N > 0 ? p.N : 1) + i];
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
    if (p.ev_count > 0 && c.tid < 8) resolve_values(p, p.ev_first, b, ginst, c.tid, c.sm + SM_ZV);
    c.sync();
    // ---- run the tape slice ----
    for (int64_t e = p.ev_first; e < p.ev_first + p.ev_count; e++) {
      const viekf_event& E = p.ev[e];
      const int kind = E.kind;
      // this event's values were resolved while the previous event ran; resolve the next event's now (8 lanes of the
      // last warp), into the other buffer: the barrier that ends this event publishes them
      c.zv = c.sm + SM_ZV + 8 * (int)((e - p.ev_first) & 1);
      if (e + 1 < p.ev_first + p.ev_count && c.tid >= G - 32 && (c.tid & 31) < 8)
        resolve_values(p, e + 1, b, ginst, c.tid & 31, c.sm + SM_ZV + 8 * (int)((e + 1 - p.ev_first) & 1));
      int result = VIEKF_MEAS_NONE;
      switch (kind) {
        case VIEKF_EV_PROP:
          ev_propagate(c, p, E);
          break;
        case VIEKF_EV_UPD:
          {
            SmemCore<G> core{c, p};
            result = ev_update(c, p, E, core);
          }
          break;
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
      if (LOCK)
        __syncthreads();
      else
        c.sync();
      if (p.results && c.tid == 0 && b >= p.res_first && b < p.res_first + p.res_count)
        p.results[(b - p.res_first) * p.n_events + e] = (uint8_t)result;
    }
    // ---- store the filter ----
    c.sync();
    latch_nan_state(c, smem_nan_scan(c));  // NaNsInTheHouse at the end of the slice
    const int len1 = c.sc[SC_LEN];
    if (c.tid == 0) {
      M.len = len1; M.next_id = c.sc[SC_NEXT]; M.started = c.sc[SC_STARTED]; M.flags = (uint32_t)c.sc[SC_FLAGS];
      M.n_gated = (uint32_t)c.sc[SC_NGATED]; M.n_fix = (uint32_t)c.sc[SC_NFIX]; M.n_nan_gain = (uint32_t)c.sc[SC_NNAN]; M.n_kf = c.sc[SC_NKF];
      M.t = *c.tcur;
      for (int k = 0; k < 6; k++) M.ub[k] = c.sm[SM_UB + k];
    }
    for (int i = c.tid; i < len1; i += G) p.ids[b * (p.N > 0 ? p.N : 1) + i] = c.ids[i];
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
  }
}

template <int G>
static cudaError_t launch_g(const KParams& p, int sm_count, cudaStream_t stream) {
  constexpr int CTA = (G >= 128 ? G : (G == 64 ? VIEKF_G64_CTA : VIEKF_SMALLG_CTA));
  constexpr int GPC = CTA / G;
  const size_t smem = smem_bytes_per_filter(p.N) * GPC;
  cudaError_t e = cudaFuncSetAttribute(viekf_tape_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(viekf_tape_kernel<G>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return e;
  int occ = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, viekf_tape_kernel<G>, CTA, smem);
  if (e != cudaSuccess) return e;
  if (occ < 1) return cudaErrorInvalidConfiguration;
  const int64_t need = (p.B + GPC - 1) / GPC;
  int64_t grid = (int64_t)sm_count * occ;  // persistent: a whole number of waves of the SM count
  if (need < grid) grid = need;
  if (grid < 1) grid = 1;
  viekf_tape_kernel<G><<<(unsigned)grid, CTA, smem, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_tape(const KParams& p, int group, int sm_count, cudaStream_t stream) {
  switch (group) {
    case GROUP_TILE: return launch_tile(p, sm_count, stream);
    case GROUP_OCTET: return launch_octet(p, sm_count, stream);
    case GROUP_FRAG: return launch_frag(p, sm_count, stream);
    case 32: return launch_g<32>(p, sm_count, stream);
    case 64: return launch_g<64>(p, sm_count, stream);
    case 128: return launch_g<128>(p, sm_count, stream);
    case 192: return launch_g<192>(p, sm_count, stream);
    case 256: return launch_g<256>(p, sm_count, stream);
    default: return cudaErrorInvalidValue;
  }
}

// ------------------------------------------------------------------------------------------------------------------
// initialisation, read-out
// ------------------------------------------------------------------------------------------------------------------
__global__ void reset_kernel(const __grid_constant__ KParams p) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= p.B) return;
  double* x = p.x + b * p.xs;
  for (int i = 0; i < 17; i++) x[i] = p.x0[i];
  double* P = p.P + b * p.ps;
  for (int i = 0; i < 136; i++) P[i] = 0.0;
  for (int i = 0; i < 16; i++) P[plo(i, i)] = p.P0[i];
  FilterMeta m;
  m.t = nan("");
  for (int i = 0; i < 3; i++) m.node_t[i] = 0;
  m.node_q[0] = 1; m.node_q[1] = m.node_q[2] = m.node_q[3] = 0;
  for (int i = 0; i < 36; i++) m.gcov[i] = 0;
  for (int i = 0; i < 6; i++) m.ub[i] = 0;
  m.started = 0; m.len = 0; m.next_id = 0; m.n_kf = 0; m.flags = 0; m.n_gated = m.n_fix = m.n_nan_gain = 0;
  p.meta[b] = m;
}
cudaError_t launch_reset(const KParams& p, cudaStream_t stream) {
  reset_kernel<<<(unsigned)((p.B + 127) / 128), 128, 0, stream>>>(p);
  return cudaGetLastError();
}

__global__ void set_x0_kernel(const __grid_constant__ KParams p, const double* x0, int broadcast) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= p.B) return;
  const double* s = broadcast ? x0 : x0 + 17 * b;
  double* x = p.x + b * p.xs;
  for (int i = 0; i < 17; i++) x[i] = s[i];
}
cudaError_t launch_set_x0(const KParams& p, const double* x0_dev, int broadcast, cudaStream_t stream) {
  set_x0_kernel<<<(unsigned)((p.B + 127) / 128), 128, 0, stream>>>(p, x0_dev, broadcast);
  return cudaGetLastError();
}

// x <- x [+] (sqrt(P0) o n), n ~ N(0,1) from Philox(seed, global instance)
__global__ void perturb_kernel(const __grid_constant__ KParams p, uint64_t seed) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= p.B) return;
  double d[16];
  for (int k = 0; k < 8; k++) {
    double nn[2];
    viekf_normal_pair(seed, (uint64_t)(p.inst_off + b), VIEKF_NOISE_EVENT_X0, k, nn);
    d[2 * k] = sqrt(p.P0[2 * k]) * nn[0];
    d[2 * k + 1] = sqrt(p.P0[2 * k + 1]) * nn[1];
  }
  double* x = p.x + b * p.xs;
  for (int k = 0; k < 6; k++) x[k] += d[k];
  stq(x + 6, otimes(ldq(x + 6), qexp(mk3(d[6], d[7], d[8]))));
  for (int k = 0; k < 7; k++) x[10 + k] += d[9 + k];
}
cudaError_t launch_perturb(const KParams& p, uint64_t seed, cudaStream_t stream) {
  perturb_kernel<<<(unsigned)((p.B + 127) / 128), 128, 0, stream>>>(p, seed);
  return cudaGetLastError();
}

// get_state: [MAX_X x count], dead entries 0
__global__ void unpack_state_kernel(const __grid_constant__ KParams p, int64_t first, int64_t count, double* out) {
  const int mx = 17 + 5 * p.N;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count * mx) return;
  const int64_t b = idx / mx;
  const int i = (int)(idx - b * mx);
  const int live = 17 + 5 * p.meta[first + b].len;
  out[idx] = (i < live) ? p.x[(first + b) * p.xs + i] : 0.0;
}
cudaError_t launch_unpack_state(const KParams& p, int64_t first, int64_t count, double* o, cudaStream_t stream) {
  const int64_t tot = count * (17 + 5 * p.N);
  unpack_state_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(p, first, count, o);
  return cudaGetLastError();
}

// get_covariance: full symmetric [MAX_DX x MAX_DX x count] column-major, dead rows/cols 0
__global__ void unpack_cov_kernel(const __grid_constant__ KParams p, int64_t first, int64_t count, double* out) {
  const int n = 16 + 3 * p.N;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count * n * n) return;
  const int64_t b = idx / ((int64_t)n * n);
  const int r = (int)(idx - b * n * n);
  const int j = r / n, i = r - j * n;
  const int live = 16 + 3 * p.meta[first + b].len;
  out[idx] = (i < live && j < live) ? p.P[(first + b) * p.ps + psym(i, j)] : 0.0;
}
cudaError_t launch_unpack_cov(const KParams& p, int64_t first, int64_t count, double* o, cudaStream_t stream) {
  const int64_t n = 16 + 3 * p.N;
  const int64_t tot = count * n * n;
  unpack_cov_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(p, first, count, o);
  return cudaGetLastError();
}

// get_covariance_diagonal: [MAX_DX x count], dead entries 0
__global__ void unpack_diag_kernel(const __grid_constant__ KParams p, int64_t first, int64_t count, double* out) {
  const int n = 16 + 3 * p.N;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count * n) return;
  const int64_t b = idx / n;
  const int i = (int)(idx - b * n);
  const int live = 16 + 3 * p.meta[first + b].len;
  out[idx] = (i < live) ? p.P[(first + b) * p.ps + plo(i, i)] : 0.0;
}
cudaError_t launch_unpack_diag(const KParams& p, int64_t first, int64_t count, double* o, cudaStream_t stream) {
  const int64_t tot = count * (16 + 3 * p.N);
  unpack_diag_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(p, first, count, o);
  return cudaGetLastError();
}

// set_imu_bias(b_g, b_a): v = [b_g(3), b_a(3)]
__global__ void set_bias_kernel(const __grid_constant__ KParams p, const double* v) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= p.B) return;
  double* x = p.x + b * p.xs;
  for (int k = 0; k < 3; k++) {
    x[13 + k] = v[k];
    x[10 + k] = v[3 + k];
  }
}
cudaError_t launch_set_bias(const KParams& p, const double* v, cudaStream_t stream) {
  set_bias_kernel<<<(unsigned)((p.B + 127) / 128), 128, 0, stream>>>(p, v);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------------------------
// FP64 FMA throughput probe (roofline denominator of the compute-bound kernels)
// ------------------------------------------------------------------------------------------------------------------
// 16 independent chains per thread, 2 CTAs x 512 threads per SM: measured 37.06 TFLOP/s on B200 = 99.6 % of 148 SMs x 64 DFMA/clk x
// 1.965 GHz (tools/microbench/dmma_peak.cu, profiles/r1_dmma_vs_dfma_microbench.txt); fewer chains / more warps measure ~8 % low
__global__ void __launch_bounds__(512) dfma_probe_kernel(double* out, int iters, double a, double b) {
  double r[16];
#pragma unroll
  for (int u = 0; u < 16; u++) r[u] = threadIdx.x + u;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 16; u++) r[u] = fma(r[u], a, b);
  }
  double s = 0;
#pragma unroll
  for (int u = 0; u < 16; u++) s += r[u];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
cudaError_t fp64_probe(int device, double* tflops) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return e;
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) return e;
  // one kernel >= 60 ms (a 5 ms kernel reads ~8 % low: it ends before the clocks have settled): warm-up, then the best of three
  const int blocks = prop.multiProcessorCount * 2, threads = 512, iters = 32768 * 16;
  double* d = nullptr;
  e = cudaMalloc(&d, sizeof(double) * blocks * threads);
  if (e != cudaSuccess) return e;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0;
  for (int rep = 0; rep < 4; rep++) {
    cudaEventRecord(e0);
    dfma_probe_kernel<<<blocks, threads>>>(d, iters, 0.999999, 1e-9);
    cudaEventRecord(e1);
    e = cudaEventSynchronize(e1);
    if (e != cudaSuccess) break;
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fl = 2.0 * 16.0 * (double)iters * blocks * threads;
    const double tf = fl / (ms * 1e-3) / 1e12;
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  *tflops = best;
  return e;
}

}  // namespace vk
