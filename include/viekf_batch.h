/*
 * viekf_batch.h -- C ABI of the B200 batch engine for the VI-EKF filter core.
 *
 * The reference (byu-magicc/VI-EKF) has no FFI layer; its boundary is the public C++ surface of
 * class vi_ekf::VIEKF (reference include/vi_ekf.h:243-337).  Every entry point below names the
 * reference method it replaces.  All arithmetic is IEEE double; all matrices crossing the ABI are
 * column-major like Eigen's default (reference include/vi_ekf.h:53-61).
 *
 * One handle = B independent filter instances with NUM_FEATURES = N (a compile-time constant in the
 * reference, include/vi_ekf.h:39-48; a creation-time constant here) resident on one CUDA device.
 * Calls on one handle must be serialised by the caller (the reference's ROS wrapper holds one mutex
 * around every VIEKF call, src/vi_ekf_ros.cpp:173-175).  Host pointers are copied during the call.
 *
 * Return value: 0 = ok, <0 = VIEKF_ERR_*.  Per-filter outcomes use the reference's meas_result_t
 * values bit for bit (include/vi_ekf.h:132-138).
 */
#ifndef VIEKF_BATCH_H
#define VIEKF_BATCH_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- index maps: reference include/vi_ekf.h:87-111 -------------------------------------------- */
enum { VIEKF_xPOS = 0, VIEKF_xVEL = 3, VIEKF_xATT = 6, VIEKF_xB_A = 10, VIEKF_xB_G = 13, VIEKF_xMU = 16, VIEKF_xZ = 17 };
enum { VIEKF_dxPOS = 0, VIEKF_dxVEL = 3, VIEKF_dxATT = 6, VIEKF_dxB_A = 9, VIEKF_dxB_G = 12, VIEKF_dxMU = 15, VIEKF_dxZ = 16 };
#define VIEKF_MAX_X(N) (17 + 5 * (N))  /* reference include/vi_ekf.h:47 */
#define VIEKF_MAX_DX(N) (16 + 3 * (N)) /* reference include/vi_ekf.h:48 */

/* measurement_type_t: reference include/vi_ekf.h:113-125 (same numeric values) */
typedef enum {
  VIEKF_ACC = 0, VIEKF_ALT = 1, VIEKF_ATT = 2, VIEKF_POS = 3, VIEKF_VEL = 4, VIEKF_QZETA = 5,
  VIEKF_FEAT = 6, VIEKF_PIXEL_VEL = 7, VIEKF_DEPTH = 8, VIEKF_INV_DEPTH = 9, VIEKF_TOTAL_MEAS = 10
} viekf_meas_type;

/* meas_result_t: reference include/vi_ekf.h:132-138 (same numeric values) */
typedef enum {
  VIEKF_MEAS_SUCCESS = 0, VIEKF_MEAS_GATED = 1, VIEKF_MEAS_NAN = 2, VIEKF_MEAS_INVALID = 3,
  VIEKF_MEAS_NEW_FEATURE = 4,
  VIEKF_MEAS_NONE = 255 /* event produced no meas_result_t (propagate, keep, ...) */
} viekf_meas_result;

enum {
  VIEKF_OK = 0, VIEKF_ERR_ARG = -1, VIEKF_ERR_CUDA = -2, VIEKF_ERR_NODEVICE = -3, VIEKF_ERR_UNSUPPORTED = -4,
  VIEKF_ERR_NOMEM = -5
};

/* Per-filter sticky flags (replace the reference's debug-only NAN_CHECK / NEGATIVE_DEPTH exits,
 * include/vi_ekf.h:29-37, src/vi_ekf/vi_ekf_error.cpp). */
enum {
  VIEKF_FLAG_NAN_STATE = 1,       /* NaNsInTheHouse() (vi_ekf_error.cpp:6-19) fired: NaN in the live x or P; evaluated at every
                                     STAT event and at the end of every launch (tape slice / per-call entry point) */
  VIEKF_FLAG_DEPTH_NAN_FIXED = 2, /* fix_depth: rho was NaN   (vi_ekf_helper.cpp:137-141) */
  VIEKF_FLAG_DEPTH_NEG_FIXED = 4, /* fix_depth: rho < 0       (vi_ekf_helper.cpp:142-148) */
  VIEKF_FLAG_DEPTH_BIG_FIXED = 8, /* fix_depth: rho > 100     (vi_ekf_helper.cpp:149-154) */
  VIEKF_FLAG_NAN_GAIN = 16,       /* NO_NANS(K)&&NO_NANS(H) failed -> update skipped (vi_ekf_meas.cpp:247) */
  VIEKF_FLAG_BACKWARD_DT = 32     /* "Trying to propagate backwards" (vi_ekf.cpp:285-289) */
};

/* Everything VIEKF::load reads from params/ekf.yaml (reference src/vi_ekf/vi_ekf.cpp:101-155) and the
 * arguments of the 17-argument VIEKF::init (vi_ekf.cpp:64-99).  Quaternions are [w,x,y,z]. */
typedef struct viekf_config {
  double x0[17];
  double P0[16];
  double Qx[16];
  double lambda[16];
  double Qu[6];
  double P0_feat[3];
  double Qx_feat[3];
  double lambda_feat[3];
  double cam_center[2];
  double focal_len[2];
  double q_b_c[4];
  double p_b_c[3];
  double q_b_u[4];
  double min_depth;
  double keyframe_overlap_threshold;
  int32_t use_drag_term;
  int32_t use_partial_update;
  int32_t use_keyframe_reset;
  int32_t reserved;
} viekf_config;

/* Fills cfg with the values of the reference's params/ekf.yaml. */
void viekf_config_default(viekf_config* cfg);
/* VIEKF(param_file) / load (vi_ekf.cpp:101-155): starts from viekf_config_default and overrides every key of the YAML file
 * that load() reads (x0, P0, P0_feat, Qx, Qx_feat, Qu, lambda, lambda_feat, cam_center, focal_len, q_b_c, p_b_c, q_b_u,
 * min_depth, keyframe_overlap_threshold, use_drag_term, use_partial_update, use_keyframe_reset).  Flat `key: value` /
 * `key: [v, v, ...]` YAML (lists may span lines, `#` comments), which is all params/ekf.yaml uses.  VIEKF_ERR_ARG on an
 * unreadable file or a list of the wrong length. */
int viekf_config_load_yaml(const char* path, viekf_config* cfg);

/* ---- in-order event tape (replaces the rewind/replay queue, vi_ekf_meas.cpp:6-194) --- */
typedef enum {
  VIEKF_EV_PROP = 0,  /* propagate_state(u, t, false)                       vi_ekf.cpp:262-318 */
  VIEKF_EV_UPD = 1,   /* add_measurement + update in order; a FEAT with an untracked id initialises the
                         feature and returns MEAS_NEW_FEATURE                vi_ekf_meas.cpp:130-147,196-278 */
  VIEKF_EV_INIT = 2,  /* init_feature(pixel, id, depth)                     vi_ekf_feat.cpp:6-47 */
  VIEKF_EV_CLEAR = 3, /* clear_feature(id)                                  vi_ekf_feat.cpp:50-73 */
  VIEKF_EV_KEEP = 4,  /* keep_only_features(ids)                            vi_ekf_feat.cpp:81-142 */
  VIEKF_EV_STAT = 5,  /* accumulate error statistics against the truth carried in z (no reference
                         counterpart; names from multirotor_sim/mclogger.h:50-52) */
  VIEKF_EV_KFRESET = 6 /* keyframe_reset()                                   vi_ekf_kfr.cpp:56-157 */
} viekf_event_kind;

typedef struct viekf_event {
  int32_t kind;   /* viekf_event_kind */
  int32_t type;   /* UPD: viekf_meas_type */
  int32_t id;     /* UPD/INIT/CLEAR: feature id (-1 none); STAT: bucket index */
  int32_t active; /* UPD: the reference's `active` flag (vi_ekf_meas.cpp:230) */
  int32_t zdim;   /* UPD: rows of z (4 for ATT/QZETA quaternions) */
  int32_t rdim;   /* UPD: rows of the residual / R */
  int32_t n_ids;  /* KEEP: number of ids */
  int32_t ids_off;/* KEEP: offset into the tape's id pool */
  int32_t noise;  /* >=0: index of the per-instance value block for this event (see viekf_tape), -1: z is shared */
  int32_t aux;    /* PROP: >= 0 = index of the physical IMU sample; all PROP events of one sample (an interval split by a
                     measurement) then draw the same generated noise; -1 = key the noise by the event index.
                     UPD: 1 = take z[k], k < zdim, from the body-frame IMU input of the filter's last propagate_state
                     (the drag-accelerometer measurement of src/vi_ekf_ros.cpp:180-184), 0 = use z,
                     2 = plain update() as replayed by handle_measurements (vi_ekf_meas.cpp:84): no add_measurement bookkeeping,
                     a FEAT whose id is not tracked is MEAS_INVALID instead of initialising the feature */
  double t;       /* PROP: absolute time of the IMU sample; others: informational */
  double z[8];    /* PROP: u = [acc(3), gyro(3)] in the IMU frame; UPD: z[0..zdim), z[4] = depth (FEAT);
                     INIT: pixel(2), z[4] = depth (NaN = unknown); STAT: true pos(3) vel(3) att(4)=10 -> see truth block */
  double R[9];    /* UPD: rdim x rdim column-major */
  double sigma[8]; /* per-instance white noise added to z[k] (PROP: k<6, UPD: k<zdim, INIT: k<2) when the tape has a
                      seed: z[k] += sigma[k] * N(0,1)(seed, instance, event index, k)   (include/viekf_noise.h) */
} viekf_event;

/* A tape is shared by all B instances (same control flow, Monte-Carlo use).  Per-instance values, when
 * present, live in `values`: instance b reads doubles values[b*values_stride + 8*noise + k], k<8, in place of
 * event.z[k] (PROP: k<6, UPD: k<zdim, STAT: 16 true base-state entries are read from two consecutive blocks). */
typedef struct viekf_tape {
  const viekf_event* events;
  int64_t n_events;
  const int32_t* id_pool;
  int64_t n_ids;
  const double* values; /* host pointer, may be NULL */
  int64_t values_stride; /* doubles per instance */
  uint64_t seed;         /* 0 = no generated noise; else Philox key for sigma / imu_bias_range (viekf_noise.h) */
  double imu_bias_range[6]; /* PROP: u[k] += imu_bias_range[k] * U(-1,1)(seed, instance, k), constant per instance
                               (multirotor_sim draws its IMU biases the same way, lib/multirotor_sim/src/simulator.cpp:140-160);
                               STAT events add the same bias (rotated by q_b_u) to the true b_a / b_g */
} viekf_tape;

/* Statistics accumulated by STAT events, per bucket (all sums over the filters of this handle):
 * [0] count, [1..16] sum e, [17..32] sum e^2, [33] sum NEES16, [34] sum NEES6 (pos+att),
 * [35] #gated updates so far, [36] #fix_depth interventions so far, [37] / [39] number of filters whose NEES16 / NEES6 was
 * defined (a block that is not positive definite -- exactly zero position covariance right after a keyframe reset -- has no
 * NEES and is left out of [33] / [34]), [38] number of filters with VIEKF_FLAG_NAN_STATE (excluded from every other column).
 * e = x_true [-] x_hat on the 16 base error states (reference boxminus, vi_ekf_helper.cpp:100-111); the global truth is first
 * expressed in the frame of the filter's current keyframe node (inverse of get_global_pose, vi_ekf_kfr.cpp:14-22), which is
 * the identity until the first keyframe_reset. */
#define VIEKF_STAT_WIDTH 40

typedef struct viekf_batch viekf_batch; /* opaque */

/* VIEKF::VIEKF() + init(...)/load(...) for B instances (vi_ekf.cpp:7-155). device = CUDA ordinal. */
int viekf_batch_create(const viekf_config* cfg, int64_t B, int32_t N, int32_t device, viekf_batch** out);
void viekf_batch_destroy(viekf_batch* h);
const char* viekf_last_error(void);

/* set_x0 / set_imu_bias generalised: overwrite the first 17 state entries of every instance
 * (vi_ekf.cpp:157-160,180-184).  x0 is [17 x B] column-major (instance b at x0 + 17*b), or one column when
 * broadcast != 0. */
int viekf_batch_set_x0(viekf_batch* h, const double* x0, int32_t broadcast);
/* Perturb every instance: x <- x [+] (sqrt(P0) .* n), n ~ N(0,1) from Philox(seed, instance). */
int viekf_batch_perturb_x0(viekf_batch* h, uint64_t seed);

/* get_state (vi_ekf.cpp:187-190): x is [MAX_X x count], instances first..first+count.  Only the live part
 * x[0 : 17+5*len) is defined; dead entries are written as 0. */
int viekf_batch_get_state(viekf_batch* h, int64_t first, int64_t count, double* x);
/* viekf_batch_get_state without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and output as viekf_batch_get_state; the device scratch between the engine's layout and the copy to `x` is
 * a fixed file-scope device array of VIEKF_NOALLOC_SCRATCH_DOUBLES doubles (2 MiB, held by every GPU that loads the library
 * until the process exits) instead of the handle's grow-only staging buffer, so the call issues no cudaMalloc / cudaFree.
 * count * MAX_X(N) must not exceed VIEKF_NOALLOC_SCRATCH_DOUBLES: VIEKF_ERR_ARG (with the limit in viekf_last_error)
 * otherwise -- read the batch in several calls.  The array is shared by every handle and every call on a device and keeps the
 * last call's contents; each call overwrites all of the part it copies out, so nothing needs clearing, but calls that overlap
 * on the GPU (two handles' streams, two host threads) race on it: serialise them. */
#define VIEKF_NOALLOC_SCRATCH_DOUBLES 262144
int viekf_batch_get_state_noalloc(viekf_batch* h, int64_t first, int64_t count, double* x);
/* get_covariance (vi_ekf.cpp:197-200): P is [MAX_DX x MAX_DX x count] column-major, reference ordering. */
int viekf_batch_get_covariance(viekf_batch* h, int64_t first, int64_t count, double* P);
/* get_len_features / tracked_features (include/vi_ekf.h:272,286): len[count], ids[N x count] (-1 padded). */
int viekf_batch_get_features(viekf_batch* h, int64_t first, int64_t count, int32_t* len, int32_t* ids);
/* sticky VIEKF_FLAG_* per filter, counters[3 x count] = {#gated, #fix_depth, #nan_gain}. */
int viekf_batch_get_flags(viekf_batch* h, int64_t first, int64_t count, uint32_t* flags, uint32_t* counters);
/* get_global_pose / global_pose_cov_ bookkeeping of keyframe_reset (vi_ekf_kfr.cpp:14-55): node pose [t(3), q(4)] x count
 * and the accumulated 6x6 global covariance (column-major, may be NULL). */
int viekf_batch_get_global_pose(viekf_batch* h, int64_t first, int64_t count, double* node_t3q4, double* gcov36);
/* get_covariance_diagonal (vi_ekf.cpp:202-206): d is [MAX_DX x count]; entries beyond the live 16 + 3 len are 0. */
int viekf_batch_get_covariance_diagonal(viekf_batch* h, int64_t first, int64_t count, double* d);
/* get_depths / get_zetas / get_qzetas (vi_ekf.cpp:210-240): depths [N x count] = 1 / rho, zetas [3 x N x count] = q_zeta.rota(e_z),
 * qzetas [4 x N x count]; entries of feature slots beyond len_features are NaN (the reference returns len_features columns). */
int viekf_batch_get_depths(viekf_batch* h, int64_t first, int64_t count, double* depths);
int viekf_batch_get_zetas(viekf_batch* h, int64_t first, int64_t count, double* zetas);
int viekf_batch_get_qzetas(viekf_batch* h, int64_t first, int64_t count, double* qzetas);
/* get_depth(id) / get_feat(id) (vi_ekf.cpp:242-260): depth[count], pixel[2 x count] of the feature with that (internal) id;
 * NaN for an instance that does not track it (the reference indexes the state with -1 there). get_zeta(i) is column i of
 * viekf_batch_get_zetas. */
int viekf_batch_get_depth(viekf_batch* h, int32_t id, int64_t first, int64_t count, double* depth);
int viekf_batch_get_feat(viekf_batch* h, int32_t id, int64_t first, int64_t count, double* pixel);
/* set_imu_bias(b_g, b_a) (vi_ekf.cpp:180-184) on every instance; set_drag_term / get_drag_term (include/vi_ekf.h:291-292). */
int viekf_batch_set_imu_bias(viekf_batch* h, const double* b_g, const double* b_a);
int viekf_batch_set_drag_term(viekf_batch* h, int32_t use_drag_term);
int viekf_batch_get_drag_term(const viekf_batch* h);
/* current filter time t_[i_] of each instance */
int viekf_batch_get_time(viekf_batch* h, int64_t first, int64_t count, double* t);

/* propagate_state(u, t, save_input) (vi_ekf.cpp:262-318).  u is [6 x B] (IMU frame), or one column when
 * broadcast != 0.  The first call only latches the start time (vi_ekf.cpp:274-279). */
int viekf_batch_propagate(viekf_batch* h, const double* u, int32_t broadcast, double t);

/* In-order update(meas) (vi_ekf_meas.cpp:196-278) at the filters' current time, preceded by
 * add_measurement's feature bookkeeping (vi_ekf_meas.cpp:136-147): z is [zdim x B] or one column when
 * broadcast != 0; R is rdim x rdim column-major; result[B] receives meas_result_t per filter (may be NULL). */
int viekf_batch_update(viekf_batch* h, int32_t type, const double* z, int32_t zdim, int32_t broadcast,
                       const double* R, int32_t rdim, int32_t active, int32_t id, double depth,
                       uint8_t* result);

/* init_feature / clear_feature / keep_only_features (vi_ekf_feat.cpp:6-142), applied to every instance.
 * pixel is [2 x B] or broadcast. ok[B] receives init_feature's bool (may be NULL). */
int viekf_batch_init_feature(viekf_batch* h, const double* pixel, int32_t broadcast, int32_t id, double depth,
                             uint8_t* ok);
int viekf_batch_clear_feature(viekf_batch* h, int32_t id);
int viekf_batch_keep_only_features(viekf_batch* h, const int32_t* ids, int32_t n);
int viekf_batch_keyframe_reset(viekf_batch* h);

/* ---- Delayed-measurement queue: the reference's streaming call pattern (vi_ekf_meas.cpp:6-194) ------------------------
 * The reference keeps a LEN_STATE_HIST-deep ring of (x, P, t) and a sorted measurement deque (include/vi_ekf.h:50-51,
 * 157-181); handle_measurements rewinds the ring to just before the oldest pending measurement and replays the stored
 * inputs and measurements.  viekf_batch_queue_enable gives the handle the same ring in HBM (len_state_hist slots of x and
 * packed P for all B instances; VIEKF_ERR_NOMEM when it does not fit) and the same two deques on the host.  Call times and
 * values are shared by all instances (the instances differ by their state, e.g. viekf_batch_perturb_x0), so the
 * rewind / replay decisions are uniform over the batch while gating stays per instance.  Must be called before the
 * first propagate.  Pass 250 / 200 for the reference's LEN_STATE_HIST / LEN_MEAS_HIST. */
int viekf_batch_queue_enable(viekf_batch* h, int32_t len_state_hist, int32_t len_meas_hist);
/* propagate_state(u, t, save_input) (vi_ekf.cpp:262-318) on the ring: the new state goes to the next slot, the input is
 * pushed onto the input deque when save_input != 0.  u is one IMU sample [6] (IMU frame). */
int viekf_batch_queue_propagate(viekf_batch* h, const double* u, double t, int32_t save_input);
/* add_measurement (vi_ekf_meas.cpp:130-194): t < start => MEAS_INVALID, NaN in z => MEAS_NAN, FEAT with an untracked id =>
 * init_feature now and MEAS_NEW_FEATURE, otherwise sorted insert (LIFO among equal times) and MEAS_SUCCESS.
 * z is [zdim], R rdim x rdim column-major; *result receives the meas_result_t. */
int viekf_batch_add_measurement(viekf_batch* h, double t, int32_t type, const double* z, int32_t zdim, const double* R,
                                int32_t rdim, int32_t active, int32_t id, double depth, int32_t* result);
/* handle_measurements (vi_ekf_meas.cpp:6-127).  For the k-th update() applied by this call: upd_types[k], upd_ids[k] and
 * upd_results[k*B + b] = meas_result_t of instance b (the reference's gated_feature_ids list is {upd_ids[k] : FEAT and
 * result == MEAS_GATED}, per instance).  cap = capacity of the three arrays in updates (any may be NULL);
 * *n_updates receives the number applied. */
int viekf_batch_handle_measurements(viekf_batch* h, int32_t* upd_types, int32_t* upd_ids, uint8_t* upd_results, int32_t cap,
                                    int32_t* n_updates);

/* Fused path: run a whole tape with x and P resident on chip.  results (may be NULL) is
 * [n_events x result_count] uint8 meas_result_t for instances result_first.. (row-major by instance:
 * results[(b-result_first)*n_events + e]). */
int viekf_batch_run_tape(viekf_batch* h, const viekf_tape* tape, uint8_t* results, int64_t result_first,
                         int64_t result_count);

/* Upload a tape once (events, ids, per-instance values) and run it later without host<->device traffic. */
typedef struct viekf_dev_tape viekf_dev_tape;
int viekf_batch_upload_tape(viekf_batch* h, const viekf_tape* tape, viekf_dev_tape** out);
void viekf_batch_free_tape(viekf_batch* h, viekf_dev_tape* t);
/* n_first..n_first+n_run events; returns after enqueueing when async != 0 (viekf_batch_sync to wait). */
int viekf_batch_run_dev_tape(viekf_batch* h, const viekf_dev_tape* t, int64_t ev_first, int64_t ev_count, int32_t async);
int viekf_batch_sync(viekf_batch* h);
/* cudaStream_t the handle launches on, as void* (for CUDA-event timing by the caller). */
void* viekf_batch_stream(viekf_batch* h);
/* Kernel launches issued by this handle so far. */
int64_t viekf_batch_launch_count(const viekf_batch* h);

/* Statistics: n_buckets x VIEKF_STAT_WIDTH doubles.  dev_ptr variant exposes the device buffer so the caller can
 * all-reduce it in place with NCCL before reading it. */
int viekf_batch_stats_reset(viekf_batch* h, int32_t n_buckets);
int viekf_batch_stats(viekf_batch* h, double* out, int32_t n_buckets);
void* viekf_batch_stats_dev_ptr(viekf_batch* h);

/* Reset every instance to the configuration's initial state (re-run benchmarks without re-creating). */
int viekf_batch_reset(viekf_batch* h);

/* Global id of this handle's instance 0 (default 0).  Generated noise is keyed by the global instance id, so a batch
 * sharded over several handles / GPUs draws exactly the streams the unsharded batch would. */
int viekf_batch_set_instance_offset(viekf_batch* h, int64_t first_global_instance);

/* FP64 FMA throughput probe (TFLOP/s, 2 flop per DFMA) of the handle's device: the roofline denominator of the
 * compute-bound kernels (MEASURED_PEAKS.json carries no FP64 figure). */
int viekf_fp64_peak_probe(int32_t device, double* tflops);

#ifdef __cplusplus
}
#endif
#endif
