"""TEST INFRASTRUCTURE -- ctypes binding of oracle/liboracle.so (CPU restatement of the reference filter core).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def build():
    subprocess.check_call(["make", "-C", _HERE, "-s"])


def lib():
    global _lib
    if _lib is None:
        p = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(p):
            build()
        L = C.CDLL(p)
        L.vo_create.restype = C.c_void_p
        L.vo_time.restype = C.c_double
        L.vo_run_batch.restype = C.c_double
        L.vo_flags.restype = C.c_uint
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleVIEKF:
    """One reference-shaped filter (reference include/vi_ekf.h:243-337) backed by the CPU restatement."""

    def __init__(self, cfg, NF, hist=250):
        self.L = lib()
        self.NF, self.MAX_X, self.MAX_DX = NF, 17 + 5 * NF, 16 + 3 * NF
        self.cfg = cfg
        self.h = C.c_void_p(self.L.vo_create(C.byref(cfg), C.c_int(NF), C.c_int(hist)))

    def __del__(self):
        try:
            self.L.vo_destroy(self.h)
        except Exception:
            pass

    def propagate_state(self, u, t, save_input=True):
        u = np.ascontiguousarray(u, dtype=np.float64)
        self.L.vo_propagate(self.h, _p(u), C.c_double(t), C.c_int(int(save_input)))

    def add_measurement(self, t, z, mtype, R, active=False, id=-1, depth=float("nan")):
        z = np.ascontiguousarray(z, dtype=np.float64).ravel()
        R = np.asfortranarray(np.atleast_2d(np.asarray(R, dtype=np.float64)))
        return self.L.vo_add_measurement(self.h, C.c_double(t), _p(z), C.c_int(z.size), C.c_int(mtype), _p(R), C.c_int(R.shape[1]),
                                         C.c_int(int(active)), C.c_int(id), C.c_double(depth))

    def handle_measurements(self):
        g = np.zeros(1024, dtype=np.int32)
        n = self.L.vo_handle_measurements(self.h, _p(g), C.c_int(g.size))
        return g[:n].tolist()

    def update_direct(self, mtype, z, R, active=True, id=-1, depth=float("nan")):
        z = np.ascontiguousarray(z, dtype=np.float64).ravel()
        R = np.asfortranarray(np.atleast_2d(np.asarray(R, dtype=np.float64)))
        return self.L.vo_update_direct(self.h, C.c_int(mtype), _p(z), C.c_int(z.size), _p(R), C.c_int(R.shape[1]), C.c_int(int(active)),
                                       C.c_int(id), C.c_double(depth))

    def init_feature(self, l, id, depth=-1.0):
        l = np.ascontiguousarray(l, dtype=np.float64)
        return bool(self.L.vo_init_feature(self.h, _p(l), C.c_int(id), C.c_double(depth)))

    def clear_feature(self, id):
        self.L.vo_clear_feature(self.h, C.c_int(id))

    def keep_only_features(self, ids):
        a = np.ascontiguousarray(ids, dtype=np.int32)
        self.L.vo_keep_only_features(self.h, _p(a), C.c_int(a.size))

    def keyframe_reset(self):
        self.L.vo_keyframe_reset(self.h)

    def fix_depth(self):
        self.L.vo_fix_depth(self.h)

    def get_state(self):
        x = np.zeros(self.MAX_X)
        self.L.vo_get_state(self.h, _p(x))
        return x

    def set_state(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        assert x.size == self.MAX_X
        self.L.vo_set_state(self.h, _p(x))

    def get_covariance(self):
        P = np.zeros((self.MAX_DX, self.MAX_DX), order="F")
        self.L.vo_get_cov(self.h, _p(P))
        return P

    def set_covariance(self, P):
        P = np.asfortranarray(P, dtype=np.float64)
        self.L.vo_set_cov(self.h, _p(P))

    def get_len_features(self):
        return self.L.vo_len_features(self.h)

    def tracked_features(self):
        a = np.zeros(max(self.NF, 1), dtype=np.int32)
        self.L.vo_tracked_features(self.h, _p(a))
        return a[: self.get_len_features()].tolist()

    def time(self):
        return self.L.vo_time(self.h)

    def flags(self):
        return self.L.vo_flags(self.h)

    def counters(self):
        c = np.zeros(3, dtype=np.uint32)
        self.L.vo_counters(self.h, _p(c))
        return c

    def margins(self):
        m = np.zeros(2)
        self.L.vo_margins(self.h, _p(m))
        return m

    def dynamics(self, x, u):
        x = np.ascontiguousarray(x, dtype=np.float64)
        u = np.ascontiguousarray(u, dtype=np.float64)
        dx = np.zeros(self.MAX_DX)
        A = np.zeros((self.MAX_DX, self.MAX_DX), order="F")
        G = np.zeros((self.MAX_DX, 6), order="F")
        self.L.vo_dynamics(self.h, _p(x), _p(u), _p(dx), _p(A), _p(G))
        return dx, A, G

    def boxplus(self, x, dx):
        x = np.ascontiguousarray(x, dtype=np.float64)
        dx = np.ascontiguousarray(dx, dtype=np.float64)
        out = np.zeros(self.MAX_X)
        self.L.vo_boxplus(self.h, _p(x), _p(dx), _p(out))
        return out

    def boxminus(self, x1, x2):
        x1 = np.ascontiguousarray(x1, dtype=np.float64)
        x2 = np.ascontiguousarray(x2, dtype=np.float64)
        out = np.zeros(self.MAX_DX)
        self.L.vo_boxminus(self.h, _p(x1), _p(x2), _p(out))
        return out

    def h_model(self, mtype, x, id=-1):
        x = np.ascontiguousarray(x, dtype=np.float64)
        zhat = np.zeros(4)
        H = np.zeros((3, self.MAX_DX), order="F")
        self.L.vo_h(self.h, C.c_int(mtype), _p(x), C.c_int(id), _p(zhat), _p(H))
        return zhat, H

    def keyframe_reset_jac(self, xm):
        xm = np.ascontiguousarray(xm, dtype=np.float64)
        xp = np.zeros(self.MAX_X)
        N = np.zeros((self.MAX_DX, self.MAX_DX), order="F")
        self.L.vo_keyframe_reset_jac(self.h, _p(xm), _p(xp), _p(N))
        return xp, N

    def global_pose(self):
        t = np.zeros(7)
        cov = np.zeros((6, 6), order="F")
        self.L.vo_global_pose(self.h, _p(t), _p(cov))
        return t, cov

    def run_tape(self, events, id_pool=None, values=None, n_buckets=0, seed=0, bias_range=None, ginst=0, ev_first=0, ev_count=None):
        events = np.ascontiguousarray(events)
        res = np.full(len(events), 255, dtype=np.uint8)
        stats = np.zeros((max(n_buckets, 1), 40))
        idp = np.ascontiguousarray(id_pool, dtype=np.int32) if id_pool is not None else None
        vals = np.ascontiguousarray(values, dtype=np.float64) if values is not None else None
        br = np.ascontiguousarray(bias_range if bias_range is not None else np.zeros(6), dtype=np.float64)
        n = len(events) - ev_first if ev_count is None else ev_count
        self.L.vo_run_tape(self.h, _p(events), C.c_int64(n), _p(idp), _p(vals), _p(res), _p(stats), C.c_int(n_buckets),
                           C.c_uint64(seed), _p(br), C.c_uint64(ginst), C.c_int64(ev_first))
        return res, stats[:n_buckets]

    def perturb_x0(self, seed, ginst=0):
        self.L.vo_perturb_x0(self.h, C.c_uint64(seed), C.c_uint64(ginst))


def run_batch(cfg, NF, events, id_pool=None, values=None, x0=None, count=1, nthreads=1, want_P=True, n_buckets=0, seed=0,
              bias_range=None, inst_off=0, perturb_seed=0, want_results=True):
    """`count` instances of one tape on `nthreads` host threads; returns dict with wall seconds and final x/P/..."""
    L = lib()
    events = np.ascontiguousarray(events)
    MAX_X, MAX_DX = 17 + 5 * NF, 16 + 3 * NF
    idp = np.ascontiguousarray(id_pool, dtype=np.int32) if id_pool is not None else None
    stride = 0
    vals = None
    if values is not None:
        vals = np.ascontiguousarray(values, dtype=np.float64)
        assert vals.shape[0] >= count
        stride = vals.shape[1]
    x0a = np.ascontiguousarray(x0, dtype=np.float64) if x0 is not None else None
    x = np.zeros((count, MAX_X))
    P = np.zeros((count, MAX_DX, MAX_DX)) if want_P else None
    res = np.zeros((count, len(events)), dtype=np.uint8) if want_results else None
    stats = np.zeros((max(n_buckets, 1), 40))
    br = np.ascontiguousarray(bias_range if bias_range is not None else np.zeros(6), dtype=np.float64)
    lens = np.zeros(count, dtype=np.int32)
    ids = np.zeros((count, max(NF, 1)), dtype=np.int32)
    flags = np.zeros((count, 4), dtype=np.uint32)
    margins = np.zeros((count, 2))
    secs = L.vo_run_batch(C.byref(cfg), C.c_int(NF), _p(events), C.c_int64(len(events)), _p(idp), _p(vals), C.c_int64(stride),
                          _p(x0a), C.c_int64(count), C.c_int(nthreads), _p(x), _p(P), _p(res), _p(stats), C.c_int(n_buckets),
                          _p(lens), _p(ids) if NF > 0 else None, _p(flags), _p(margins), C.c_uint64(seed), _p(br), C.c_int64(inst_off),
                          C.c_uint64(perturb_seed))
    if P is not None:
        P = P.transpose(0, 2, 1)  # column-major per instance -> P[b, i, j]
    return dict(secs=secs, x=x, P=P, results=res, stats=stats[:n_buckets], len=lens, ids=ids[:, :NF], flags=flags, margins=margins)


def quat_op(op, a, b=None, n_out=4):
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b if b is not None else np.zeros(4), dtype=np.float64)
    out = np.zeros(9)
    lib().vo_quat_op(C.c_int(op), _p(a), _p(b), _p(out))
    return out[:n_out]
