"""BatchVIEKF: B reference-shaped filters behind the C ABI (include/viekf_batch.h).

Method names, argument meaning and result codes follow the reference's vi_ekf::VIEKF
(reference include/vi_ekf.h:243-337); every call acts on all B instances.
"""
import ctypes as C

import numpy as np

from . import abi


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class ViekfError(RuntimeError):
    pass


class BatchVIEKF:
    def __init__(self, cfg=None, B=1, N=12, device=0):
        self.L = abi.load_library()
        self.cfg = cfg if cfg is not None else abi.default_config()
        self.B, self.N = int(B), int(N)
        self.MAX_X, self.MAX_DX = abi.max_x(N), abi.max_dx(N)
        h = C.c_void_p()
        self._chk(self.L.viekf_batch_create(C.byref(self.cfg), self.B, self.N, int(device), C.byref(h)))
        self.h = h

    def _chk(self, rc):
        if rc != 0:
            raise ViekfError(f"viekf_b200 error {rc}: {self.L.viekf_last_error().decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.L.viekf_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- state access (reference vi_ekf.cpp:157-206) -------------------------------------------------
    def set_x0(self, x0):
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        if x0.ndim == 1:
            assert x0.size == 17
            self._chk(self.L.viekf_batch_set_x0(self.h, _p(x0), 1))
        else:
            assert x0.shape == (self.B, 17)
            self._chk(self.L.viekf_batch_set_x0(self.h, _p(x0), 0))

    def perturb_x0(self, seed):
        self._chk(self.L.viekf_batch_perturb_x0(self.h, C.c_uint64(seed)))

    def get_state(self, first=0, count=None):
        count = self.B - first if count is None else count
        x = np.zeros((count, self.MAX_X))
        self._chk(self.L.viekf_batch_get_state(self.h, first, count, _p(x)))
        return x

    def get_state_noalloc(self, first=0, count=None):
        """get_state through the fixed device scratch array: no cudaMalloc (operators' addition, include/viekf_batch.h)."""
        count = self.B - first if count is None else count
        x = np.zeros((count, self.MAX_X))
        self._chk(self.L.viekf_batch_get_state_noalloc(self.h, first, count, _p(x)))
        return x

    def get_covariance(self, first=0, count=None):
        count = self.B - first if count is None else count
        P = np.zeros((count, self.MAX_DX, self.MAX_DX))
        self._chk(self.L.viekf_batch_get_covariance(self.h, first, count, _p(P)))
        return P.transpose(0, 2, 1)  # column-major per instance -> P[b, i, j]

    def get_features(self, first=0, count=None):
        count = self.B - first if count is None else count
        ln = np.zeros(count, dtype=np.int32)
        ids = np.full((count, max(self.N, 1)), -1, dtype=np.int32)
        self._chk(self.L.viekf_batch_get_features(self.h, first, count, _p(ln), _p(ids)))
        return ln, ids[:, : self.N]

    def get_flags(self, first=0, count=None):
        count = self.B - first if count is None else count
        fl = np.zeros(count, dtype=np.uint32)
        ctr = np.zeros((count, 3), dtype=np.uint32)
        self._chk(self.L.viekf_batch_get_flags(self.h, first, count, _p(fl), _p(ctr)))
        return fl, ctr

    def get_global_pose(self, first=0, count=None):
        count = self.B - first if count is None else count
        pose = np.zeros((count, 7))
        cov = np.zeros((count, 6, 6))
        self._chk(self.L.viekf_batch_get_global_pose(self.h, first, count, _p(pose), _p(cov)))
        return pose, cov.transpose(0, 2, 1)

    # ---- feature / covariance read-back (reference vi_ekf.cpp:202-260) ------------------------------------
    def get_covariance_diagonal(self, first=0, count=None):
        count = self.B - first if count is None else count
        d = np.zeros((count, self.MAX_DX))
        self._chk(self.L.viekf_batch_get_covariance_diagonal(self.h, first, count, _p(d)))
        return d

    def get_depths(self, first=0, count=None):
        """1 / rho per feature slot, NaN beyond len_features."""
        count = self.B - first if count is None else count
        d = np.full((count, max(self.N, 1)), np.nan)
        self._chk(self.L.viekf_batch_get_depths(self.h, first, count, _p(d)))
        return d[:, : self.N]

    def get_zetas(self, first=0, count=None):
        count = self.B - first if count is None else count
        z = np.full((count, max(self.N, 1), 3), np.nan)
        self._chk(self.L.viekf_batch_get_zetas(self.h, first, count, _p(z)))
        return z[:, : self.N]

    def get_qzetas(self, first=0, count=None):
        count = self.B - first if count is None else count
        q = np.full((count, max(self.N, 1), 4), np.nan)
        self._chk(self.L.viekf_batch_get_qzetas(self.h, first, count, _p(q)))
        return q[:, : self.N]

    def get_zeta(self, i, first=0, count=None):
        return self.get_zetas(first, count)[:, i]

    def get_depth(self, id, first=0, count=None):
        count = self.B - first if count is None else count
        d = np.zeros(count)
        self._chk(self.L.viekf_batch_get_depth(self.h, int(id), first, count, _p(d)))
        return d

    def get_feat(self, id, first=0, count=None):
        count = self.B - first if count is None else count
        px = np.zeros((count, 2))
        self._chk(self.L.viekf_batch_get_feat(self.h, int(id), first, count, _p(px)))
        return px

    def set_imu_bias(self, b_g, b_a):
        bg = np.ascontiguousarray(b_g, dtype=np.float64)
        ba = np.ascontiguousarray(b_a, dtype=np.float64)
        assert bg.shape == (3,) and ba.shape == (3,)
        self._chk(self.L.viekf_batch_set_imu_bias(self.h, _p(bg), _p(ba)))

    def set_drag_term(self, use_drag_term):
        self._chk(self.L.viekf_batch_set_drag_term(self.h, int(bool(use_drag_term))))

    def get_drag_term(self):
        return bool(self.L.viekf_batch_get_drag_term(self.h))

    def set_instance_offset(self, first_global_instance):
        self._chk(self.L.viekf_batch_set_instance_offset(self.h, int(first_global_instance)))

    def get_time(self, first=0, count=None):
        count = self.B - first if count is None else count
        t = np.zeros(count)
        self._chk(self.L.viekf_batch_get_time(self.h, first, count, _p(t)))
        return t

    # ---- per-call path (one launch per reference call) -----------------------------------------------
    def propagate_state(self, u, t):
        u = np.ascontiguousarray(u, dtype=np.float64)
        bc = 1 if u.ndim == 1 else 0
        assert u.shape[-1] == 6 and (bc or u.shape[0] == self.B)
        self._chk(self.L.viekf_batch_propagate(self.h, _p(u), bc, float(t)))

    def update(self, mtype, z, R, active=True, id=-1, depth=float("nan")):
        z = np.ascontiguousarray(z, dtype=np.float64)
        R = np.asfortranarray(np.atleast_2d(np.asarray(R, dtype=np.float64)))
        bc = 1 if z.ndim == 1 else 0
        zdim = z.shape[-1]
        res = np.zeros(self.B, dtype=np.uint8)
        self._chk(self.L.viekf_batch_update(self.h, int(mtype), _p(z), zdim, bc, _p(R), R.shape[0], int(active), int(id),
                                            float(depth), _p(res)))
        return res

    def init_feature(self, pixel, id, depth=float("nan")):  # NaN = unknown depth, 2 * min_depth (vi_ekf_feat.cpp:24-28)
        pixel = np.ascontiguousarray(pixel, dtype=np.float64)
        bc = 1 if pixel.ndim == 1 else 0
        ok = np.zeros(self.B, dtype=np.uint8)
        self._chk(self.L.viekf_batch_init_feature(self.h, _p(pixel), bc, int(id), float(depth), _p(ok)))
        return ok.astype(bool)

    def clear_feature(self, id):
        self._chk(self.L.viekf_batch_clear_feature(self.h, int(id)))

    def keep_only_features(self, ids):
        a = np.ascontiguousarray(ids, dtype=np.int32)
        self._chk(self.L.viekf_batch_keep_only_features(self.h, _p(a), a.size))

    def keyframe_reset(self):
        self._chk(self.L.viekf_batch_keyframe_reset(self.h))

    # ---- streaming path: the reference's delayed-measurement queue (vi_ekf_meas.cpp:6-194) --------------
    def enable_queue(self, len_state_hist=250, len_meas_hist=200):
        """History ring + measurement deque (include/vi_ekf.h:50-51); afterwards propagate_state(save_input=...),
        add_measurement and handle_measurements follow the reference's rewind/replay semantics."""
        self._chk(self.L.viekf_batch_queue_enable(self.h, int(len_state_hist), int(len_meas_hist)))
        self._queue = True

    def queue_propagate_state(self, u, t, save_input=True):
        u = np.ascontiguousarray(u, dtype=np.float64)
        assert u.shape == (6,)
        self._chk(self.L.viekf_batch_queue_propagate(self.h, _p(u), float(t), int(bool(save_input))))

    def add_measurement(self, t, z, mtype, R, active=False, id=-1, depth=float("nan")):
        """Returns the meas_result_t of add_measurement (uniform over the batch)."""
        z = np.ascontiguousarray(z, dtype=np.float64).ravel()
        R = np.asfortranarray(np.atleast_2d(np.asarray(R, dtype=np.float64)))
        res = C.c_int32(-1)
        self._chk(self.L.viekf_batch_add_measurement(self.h, float(t), int(mtype), _p(z), z.size, _p(R), R.shape[0], int(bool(active)),
                                                     int(id), float(depth), C.byref(res)))
        return res.value

    def handle_measurements(self, cap=256):
        """Returns (types[k], ids[k], results[k, B]) of the updates applied by this call."""
        ty = np.zeros(cap, dtype=np.int32)
        ids = np.zeros(cap, dtype=np.int32)
        res = np.zeros((cap, self.B), dtype=np.uint8)
        n = C.c_int32(0)
        self._chk(self.L.viekf_batch_handle_measurements(self.h, _p(ty), _p(ids), _p(res), cap, C.byref(n)))
        k = min(n.value, cap)
        return ty[:k], ids[:k], res[:k]

    # ---- fused path -----------------------------------------------------------------------------------
    def _mk_tape(self, events, id_pool, values, seed=0, bias_range=None):
        events = np.ascontiguousarray(events, dtype=abi.EVENT_DTYPE)
        idp = np.ascontiguousarray(id_pool if id_pool is not None else [0], dtype=np.int32)
        t = abi.Tape()
        t.events = events.ctypes.data
        t.n_events = len(events)
        t.id_pool = idp.ctypes.data
        t.n_ids = idp.size
        t.seed = int(seed)
        if bias_range is not None:
            t.imu_bias_range[:] = [float(v) for v in bias_range]
        keep = [events, idp]
        if values is not None:
            values = np.ascontiguousarray(values, dtype=np.float64)
            assert values.ndim == 2 and values.shape[0] == self.B
            t.values = values.ctypes.data
            t.values_stride = values.shape[1]
            keep.append(values)
        return t, keep

    def run_tape(self, events, id_pool=None, values=None, results_first=0, results_count=0, seed=0, bias_range=None):
        """Whole tape through the C ABI with HOST buffers: upload, one fused launch, results back."""
        t, keep = self._mk_tape(events, id_pool, values, seed, bias_range)
        res = np.zeros((results_count, len(events)), dtype=np.uint8) if results_count else None
        self._chk(self.L.viekf_batch_run_tape(self.h, C.byref(t), _p(res), results_first, results_count))
        return res

    def upload_tape(self, events, id_pool=None, values=None, seed=0, bias_range=None):
        t, keep = self._mk_tape(events, id_pool, values, seed, bias_range)
        d = C.c_void_p()
        self._chk(self.L.viekf_batch_upload_tape(self.h, C.byref(t), C.byref(d)))
        return d

    def free_tape(self, d):
        self.L.viekf_batch_free_tape(self.h, d)

    def run_dev_tape(self, d, first=0, count=-1, sync=True):
        self._chk(self.L.viekf_batch_run_dev_tape(self.h, d, first, count, 0 if sync else 1))

    def sync(self):
        self._chk(self.L.viekf_batch_sync(self.h))

    def stream(self):
        return self.L.viekf_batch_stream(self.h)

    def launch_count(self):
        return self.L.viekf_batch_launch_count(self.h)

    def stats_reset(self, n_buckets):
        self._chk(self.L.viekf_batch_stats_reset(self.h, n_buckets))

    def stats(self, n_buckets):
        out = np.zeros((n_buckets, abi.STAT_WIDTH))
        self._chk(self.L.viekf_batch_stats(self.h, _p(out), n_buckets))
        return out

    def stats_dev_ptr(self):
        return self.L.viekf_batch_stats_dev_ptr(self.h)

    def reset(self):
        self._chk(self.L.viekf_batch_reset(self.h))


def fp64_peak_tflops(device=0):
    """Measured FP64 FMA throughput of the device (the roofline denominator of the compute-bound kernels)."""
    L = abi.load_library()
    v = C.c_double(0.0)
    rc = L.viekf_fp64_peak_probe(int(device), C.byref(v))
    if rc != 0:
        raise ViekfError(f"viekf_fp64_peak_probe: {rc}: {L.viekf_last_error().decode()}")
    return v.value
