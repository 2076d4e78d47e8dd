// viekf_engine.h -- internal interface between the C ABI (viekf_capi.cu) and the tape-executor kernels
// (viekf_kernel.cu).  Not part of the public boundary (that is include/viekf_batch.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/viekf_batch.h"

namespace vk {

// Per-filter bookkeeping that the reference keeps in VIEKF members (include/vi_ekf.h:157-239): filter time,
// feature counters, sticky error flags, and the keyframe-reset global pose / covariance (vi_ekf_kfr.cpp).
struct FilterMeta {
  double t;          // t_[i_]
  double node_t[3];  // current_node_global_pose_.t
  double node_q[4];  // current_node_global_pose_.q
  double gcov[36];   // global_pose_cov_ (column-major 6x6)
  double ub[6];      // body-frame IMU input of the last propagate_state call
  int32_t started;   // !isnan(start_t_)
  int32_t len;       // len_features_
  int32_t next_id;   // next_feature_id_
  int32_t n_kf;      // keyframe_features_.size()
  uint32_t flags;    // sticky VIEKF_FLAG_*
  uint32_t n_gated, n_fix, n_nan_gain;
};

static const int KF_CAP = 64;  // capacity of the stored keyframe feature list per filter

struct KParams {
  // ---- configuration (viekf_config, plus constants derived from it on the host) ----
  double x0[17], P0[16], Qx[16], lam[16], Qu[6], P0f[3], Qxf[3], lamf[3];
  double cc[2], fl[2], qbc[4], pbc[3], qbu[4];
  double Rbc[9];    // q_b_c.R(), row-major
  double RbcSp[9];  // q_b_c.R() * skew(p_b_c)
  double min_depth, kf_thresh;
  double depth_reset;  // 1 / (2 min_depth): the inverse depth fix_depth resets to (vi_ekf_helper.cpp:131), precomputed on the host
  int32_t use_drag, use_partial, use_kfreset;
  // ---- geometry of the batch ----
  int32_t N;      // NUM_FEATURES (capacity)
  int32_t xs;     // doubles per filter in x (>= 17+5N, even)
  int32_t ps;     // doubles per filter in P (packed lower triangle capacity, even)
  int64_t B;
  int64_t inst_off;
  // ---- state in HBM ----
  double* x;       // [B][xs]
  double* P;       // [B][ps]  packed lower triangle, row-major: P(i,j), i>=j at i(i+1)/2+j
  FilterMeta* meta;
  int32_t* ids;    // [B][max(N,1)]  current_feature_ids_
  int32_t* kf_ids; // [B][KF_CAP]    keyframe_features_
  // ---- tape ----
  const viekf_event* ev;
  int64_t ev_first, ev_count, n_events;
  const int32_t* id_pool;
  const double* values;
  int64_t vstride;
  uint64_t seed;
  double bias_range[6];
  // ---- outputs ----
  uint8_t* results;  // [res_count][n_events] or null
  int64_t res_first, res_count;
  double* stats;  // [n_buckets][VIEKF_STAT_WIDTH] or null
  int32_t n_buckets;
};

// threads cooperating on one filter, chosen from N; GROUP_TILE selects the register-tile executor (viekf_tile.cu)
static const int GROUP_TILE = 1;
// GROUP_OCTET selects the N = 0 executor with eight lanes per filter and the covariance in registers (viekf_octet.cu)
static const int GROUP_OCTET = 8;
// GROUP_FRAG selects the 1 <= N <= 8 executor with the covariance in FP64 tensor-core fragments, one warp per filter (viekf_frag.cu)
static const int GROUP_FRAG = 2;
size_t frag_smem_bytes(int N);
cudaError_t launch_frag(const KParams& p, int sm_count, cudaStream_t stream);
size_t octet_smem_bytes();
cudaError_t launch_octet(const KParams& p, int sm_count, cudaStream_t stream);
int group_size_for(int N);
// register-tile executor: tile warps for N (0 = N outside its range), shared memory per filter, launch
int tile_warps_for(int N);
size_t tile_smem_bytes(int N);
cudaError_t launch_tile(const KParams& p, int sm_count, cudaStream_t stream);
// filters that share one CTA of the shared-memory executor at this group size
int smem_filters_per_cta(int group);
// dynamic shared memory (bytes) one filter needs
size_t smem_bytes_per_filter(int N);
// launches the tape executor; grid sized as a multiple of the SM count. Returns cudaError_t.
cudaError_t launch_tape(const KParams& p, int group, int sm_count, cudaStream_t stream);
// initialise / reset all filters from the configuration
cudaError_t launch_reset(const KParams& p, cudaStream_t stream);
cudaError_t launch_set_x0(const KParams& p, const double* x0_dev, int broadcast, cudaStream_t stream);
cudaError_t launch_perturb(const KParams& p, uint64_t seed, cudaStream_t stream);
// read-out: unpack to the reference's layouts
cudaError_t launch_unpack_state(const KParams& p, int64_t first, int64_t count, double* x_out_dev, cudaStream_t stream);
cudaError_t launch_unpack_cov(const KParams& p, int64_t first, int64_t count, double* P_out_dev, cudaStream_t stream);
cudaError_t launch_unpack_diag(const KParams& p, int64_t first, int64_t count, double* d_out_dev, cudaStream_t stream);
cudaError_t launch_set_bias(const KParams& p, const double* bg3_ba3_dev, cudaStream_t stream);
// FP64 FMA probe
cudaError_t fp64_probe(int device, double* tflops);

}  // namespace vk
