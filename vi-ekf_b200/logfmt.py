"""Writer for the reference's raw-double log streams (log-format compatibility).

The reference logs one chosen filter to `<root><name>_*.log` files of packed little-endian doubles
(reference src/vi_ekf/vi_ekf_log.cpp:6-33 log_state, :79-117 init_logger) which matlab/plot_ekf.m:19-27 reads back with
`reshape(fread(f, 'double'), rows, [])`.  This module writes the same files for ONE instance of a batch (or for any object with
the same read API, e.g. the CPU oracle in the tests), so the reference's plotting scripts work on a batch run unchanged:

  <name>_state.log        [t, x(MAX_X)]                 1 + 17 + 5 N doubles per record
  <name>_cov.log          [t, diag P (MAX_DX)]          1 + 16 + 3 N
  <name>_input.log        [t, u(6)]                     7      (the body-frame IMU input passed by the caller)
  <name>_feat_id.log      [t, ids(N) as double, -1 pad] 1 + N
  <name>_global_pose.log  [t, pos(3), quat(4)]          8      (node pose (x) relative pose, vi_ekf_kfr.cpp:14-22)
  <name>_config.txt       the text block init_logger writes (vi_ekf_log.cpp:98-116), one `key: values` line each

Not written: <name>_xdot.log (the engine does not expose dx_), the per-measurement streams (log_measurement) and _kf / _debug.
"""
import os

import numpy as np


def _rota(q, v):
    b = np.asarray(q[1:])
    t = 2.0 * np.cross(v, b)
    return v - q[0] * t + np.cross(t, b)


def _otimes(a, b):
    return np.array([a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                     a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


class BatchInstanceView:
    """Read API of one instance of a BatchVIEKF, as the logger needs it."""

    def __init__(self, batch, instance=0):
        self.bk, self.i, self.N, self.cfg = batch, int(instance), batch.N, batch.cfg

    def state(self):
        return self.bk.get_state(self.i, 1)[0]

    def cov_diag(self):
        return self.bk.get_covariance_diagonal(self.i, 1)[0]

    def feature_ids(self):
        ln, ids = self.bk.get_features(self.i, 1)
        return [int(v) for v in ids[0][: ln[0]]]

    def node_pose(self):
        pose, _ = self.bk.get_global_pose(self.i, 1)
        return pose[0, :3], pose[0, 3:]


class StateLogger:
    STREAMS = ("state", "cov", "input", "feat_id", "global_pose")

    def __init__(self, root, name, view):
        self.view, self.N = view, view.N
        self.base = os.path.join(root, name)
        self.f = {s: open(f"{self.base}_{s}.log", "wb") for s in self.STREAMS}
        self._write_config(view.cfg)

    def _write_config(self, c):
        def row(v):
            return " ".join(repr(float(x)) for x in v)
        with open(f"{self.base}_config.txt", "w") as f:  # the line order plot_ekf.m:4-17 reads
            f.write(f"Test Num: {self.base}\n")
            f.write(f"x0 {row(c.x0[:])}\n")
            f.write(f"P0: {row(c.P0[:])}\n")
            f.write(f"P0_feat: {row(c.P0_feat[:])}\n")
            f.write(f"Qx: {row(c.Qx[:])}\n")
            f.write(f"Qx_feat: {row(c.Qx_feat[:])}\n")
            f.write(f"Qu: {row(c.Qu[:])}\n")
            f.write(f"q_b_c: {row(c.q_b_c[:])}\n")
            f.write(f"p_b_c: {row(c.p_b_c[:])}\n")
            f.write(f"lambda: {row(c.lambda_[:])}\n")
            f.write(f"lambda_feat: {row(c.lambda_feat[:])}\n")
            f.write(f"using partial_update: {int(c.use_partial_update)}\n")
            f.write(f"using keyframe reset: {int(c.use_keyframe_reset)}\n")
            f.write(f"using drag Term: {int(c.use_drag_term)}\n")
            f.write(f"keyframe overlap: {float(c.keyframe_overlap_threshold)!r}\n")
            f.write(f"num features: {self.N}\n")
            f.write(f"min_depth: {float(c.min_depth)!r}\n")

    def log_state(self, t, u_body):
        """One record per stream, as VIEKF::log_state does at the end of propagate_state (vi_ekf.cpp:316-317)."""
        x = np.asarray(self.view.state(), dtype="<f8")
        d = np.asarray(self.view.cov_diag(), dtype="<f8")
        ids = self.view.feature_ids()
        nt, nq = self.view.node_pose()
        gp = np.concatenate([nt + _rota(nq, x[0:3]), _otimes(nq, x[6:10])])
        idd = np.full(self.N, -1.0)
        idd[: len(ids)] = ids
        tt = np.array([t], dtype="<f8")
        for s, payload in (("state", x), ("cov", d), ("input", np.asarray(u_body, dtype="<f8")), ("feat_id", idd), ("global_pose", gp)):
            self.f[s].write(tt.tobytes())
            self.f[s].write(np.asarray(payload, dtype="<f8").tobytes())

    def close(self):
        for f in self.f.values():
            f.close()


def read_stream(path, rows):
    """plot_ekf.m's reader: reshape(fread(f, 'double'), rows, []) -> array [rows, records]."""
    a = np.fromfile(path, dtype="<f8")
    assert a.size % rows == 0, f"{path}: {a.size} doubles is not a whole number of {rows}-double records"
    return a.reshape(-1, rows).T
