"""Log-format compatibility: the raw-double streams of reference src/vi_ekf/vi_ekf_log.cpp:6-33,79-117
written for one filter and read back with the layout matlab/plot_ekf.m:19-27 uses (reshape(fread(f,'double'), rows, []))."""
import numpy as np
import pytest

from tests import tapes
from tests.common import abi


class _OracleView:
    """The logger's read API over the CPU oracle (the GPU test uses logfmt.BatchInstanceView)."""

    def __init__(self, f, cfg, N):
        self.f, self.cfg, self.N = f, cfg, N

    def state(self):
        x = self.f.get_state().copy()
        x[17 + 5 * self.f.get_len_features():] = 0.0
        return x

    def cov_diag(self):
        d = np.diag(self.f.get_covariance()).copy()
        d[16 + 3 * self.f.get_len_features():] = 0.0
        return d

    def feature_ids(self):
        return self.f.tracked_features()

    def node_pose(self):
        from oracle import np_oracle
        t7, _ = self.f.global_pose()  # global pose = node (x) rel: recover the node
        x = self.f.get_state()
        nq = np_oracle.q_otimes(t7[3:], np_oracle.q_inv(x[6:10]))
        return t7[:3] - np_oracle.q_rota(nq, x[0:3]), nq


def _drive(logger, step, n=25):
    for k in range(1, n + 1):
        t = 0.005 * k
        u = tapes.imu_at(t)
        step(u, t, k)
        logger.log_state(t, u)
    logger.close()


def _check_files(pkg, tmp_path, name, N, n_rec):
    lf = pkg.logfmt
    st = lf.read_stream(tmp_path / f"{name}_state.log", 1 + 17 + 5 * N)
    cv = lf.read_stream(tmp_path / f"{name}_cov.log", 1 + 16 + 3 * N)
    gp = lf.read_stream(tmp_path / f"{name}_global_pose.log", 8)
    im = lf.read_stream(tmp_path / f"{name}_input.log", 7)
    fi = lf.read_stream(tmp_path / f"{name}_feat_id.log", 1 + N)
    for a in (st, cv, gp, im, fi):
        assert a.shape[1] == n_rec and np.allclose(a[0], 0.005 * np.arange(1, n_rec + 1))
    cfg_lines = open(tmp_path / f"{name}_config.txt").read().splitlines()
    assert len(cfg_lines) == 17 and cfg_lines[0].startswith("Test Num:") and cfg_lines[15] == f"num features: {N}"
    assert cfg_lines[1].startswith("x0 ") and len(cfg_lines[1].split()) == 18 and len(cfg_lines[2].split(":")[1].split()) == 16
    return st, cv, gp, im, fi


def test_log_streams_of_the_oracle(pkg, oracle_mod, tmp_path):
    cfg = abi.default_config()
    N = 4
    f = oracle_mod.OracleVIEKF(cfg, N, hist=2)
    f.propagate_state(tapes.imu_at(0.0), 0.0, False)
    assert f.init_feature(np.array([300.0, 200.0]), 0, 5.0) and f.init_feature(np.array([350.0, 260.0]), 1, 4.0)
    log = pkg.logfmt.StateLogger(str(tmp_path), "ekf1", _OracleView(f, cfg, N))

    def step(u, t, k):
        f.propagate_state(u, t, False)
        if k == 12:
            f.keyframe_reset()
    _drive(log, step)
    st, cv, gp, im, fi = _check_files(pkg, tmp_path, "ekf1", N, 25)
    x = f.get_state()
    assert np.allclose(st[1:1 + 27, -1], x[:27]) and (st[1 + 27:, -1] == 0).all()
    assert np.allclose(cv[1:1 + 22, -1], np.diag(f.get_covariance())[:22])
    assert list(fi[1:, -1]) == [0.0, 1.0, -1.0, -1.0]
    t7, _ = f.global_pose()
    assert np.allclose(gp[1:, -1], t7, atol=1e-12)
    assert np.abs(gp[1:4, -1] - x[0:3]).max() > 0.05, "after the keyframe reset the global pose differs from the relative state"
    assert np.allclose(im[1:, -1], tapes.imu_at(0.125))


@pytest.mark.gpu
def test_log_streams_of_a_batch_instance(pkg, oracle_mod, tmp_path):
    cfg = abi.default_config()
    N, B = 4, 3
    bk = pkg.BatchVIEKF(cfg, B=B, N=N)
    f = oracle_mod.OracleVIEKF(cfg, N, hist=2)
    for o in (bk, f):
        if o is bk:
            o.propagate_state(tapes.imu_at(0.0), 0.0)
            o.init_feature(np.array([300.0, 200.0]), 0, depth=5.0)
        else:
            o.propagate_state(tapes.imu_at(0.0), 0.0, False)
            o.init_feature(np.array([300.0, 200.0]), 0, 5.0)
    log = pkg.logfmt.StateLogger(str(tmp_path), "gpu", pkg.logfmt.BatchInstanceView(bk, 1))

    def step(u, t, k):
        bk.propagate_state(u, t)
        f.propagate_state(u, t, False)
        if k == 12:
            bk.keyframe_reset()
            f.keyframe_reset()
    _drive(log, step)
    st, cv, gp, im, fi = _check_files(pkg, tmp_path, "gpu", N, 25)
    x = f.get_state()
    assert np.allclose(st[1:1 + 22, -1], x[:22], rtol=1e-9, atol=1e-12)
    assert np.allclose(cv[1:1 + 19, -1], np.diag(f.get_covariance())[:19], rtol=1e-9, atol=1e-15)
    t7, _ = f.global_pose()
    assert np.allclose(gp[1:, -1], t7, rtol=1e-9, atol=1e-12)
    assert list(fi[1:, -1]) == [0.0, -1.0, -1.0, -1.0]
