"""Streaming call pattern through the C ABI: the engine's delayed-measurement queue (history ring in HBM, rewind / replay,
viekf_batch_queue_enable / _queue_propagate / _add_measurement / _handle_measurements) against the oracle's literal
restatement of vi_ekf_meas.cpp:6-194 driven call for call with the same inputs.

Tolerance: add_measurement results, applied-update order, gate results, feature ids bit-exact; x, P within 1e-9 relative.

Dead feature slots: the reference's dense P carries P0_feat (+ Qx_feat per propagate) on the blocks of features that are not
live (vi_ekf.cpp:74-80,137-143,304) and a rewind behind an init_feature brings such a block back to life
(test_queue_new_features_behind_a_rewind); the engine keeps those rows in queue mode (csrc/viekf_capi.cu queue_dead_diag).
"""
import numpy as np
import pytest

from tests.common import REL_TOL, abi
from tests import tapes

pytestmark = pytest.mark.gpu


def _drive(pkg, oracle_mod, NF, n_imu, frame_every, delay, late_alt=False, hist=250, B=3, seed=3, drop_every=0, check_each=False, replace=False, Qxf=0.0):
    """IMU at 250 Hz; a camera frame every `frame_every` samples stamped `delay` seconds in the past (so it lands inside an
    earlier IMU interval and forces a rewind); ACC every sample; optionally an ALT measurement that arrives two IMU samples
    late.  Both filters receive exactly the same calls, in the order vi_ekf_test / the ROS wrapper issue them."""
    rng = np.random.default_rng(seed)
    cfg = abi.default_config()
    cfg.use_keyframe_reset = 0
    cfg.Qx_feat[:] = [Qxf, Qxf, 0.1 * Qxf]
    next_new = [NF]
    f = oracle_mod.OracleVIEKF(cfg, NF, hist=hist)
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.enable_queue(hist, 200)
    feat_R = np.diag([0.25, 0.25])
    acc_R = np.diag([1.0, 1.0])
    alt_R = np.array([[0.01]])
    pix = {i: np.array([80.0 + 37.0 * i, 60.0 + 29.0 * (i % 5)]) for i in range(NF)}
    applied = []
    t = 0.0
    for k in range(n_imu):
        t = 0.004 * (k + 1)
        u = np.array([0.1 * np.sin(0.05 * k), 0.05, -9.80665, 0.02 * np.cos(0.03 * k), -0.01, 0.015]) + 0.01 * rng.standard_normal(6)
        f.propagate_state(u, t, True)
        bk.queue_propagate_state(u, t, True)
        z_acc = u[:2] + 0.01 * rng.standard_normal(2)
        if k > 0:  # a measurement stamped at the start time has no older input and would block the queue for good (vi_ekf_meas.cpp:31-43)
            assert f.add_measurement(t, z_acc, abi.ACC, acc_R, True) == bk.add_measurement(t, z_acc, abi.ACC, acc_R, True)
        if late_alt and k % 10 == 9:
            z_alt = np.array([2.0 + 0.01 * rng.standard_normal()])
            ta = t - 0.0075  # older than measurements already handled: the reference's lossy corner
            assert f.add_measurement(ta, z_alt, abi.ALT, alt_R, True) == bk.add_measurement(ta, z_alt, abi.ALT, alt_R, True)
        if k == 0 or (k % frame_every == frame_every - 1 and t - delay > 0.0045):
            # k == 0: every id is new and is initialised in slot 0, which no later rewind goes behind (see module docstring)
            tc = t if k == 0 else t - delay
            if drop_every and k > 0 and (k // frame_every) % drop_every == drop_every - 1 and len(pix) > 2:
                keep = f.tracked_features()[1:]  # the tracker lost the oldest feature
                f.keep_only_features(keep)
                bk.keep_only_features(keep)
                for d in [i for i in pix if i not in keep]:
                    del pix[d]
                if replace:  # the tracker found a replacement: a new id in this frame (the filter numbers it next_feature_id_)
                    pix[next_new[0]] = np.array([100.0 + 11.0 * (next_new[0] % 40), 90.0 + 7.0 * (next_new[0] % 30)])
                    next_new[0] += 1
            for i in list(pix):
                pix[i] = pix[i] + 0.3 * rng.standard_normal(2)
                r_o = f.add_measurement(tc, pix[i], abi.FEAT, feat_R, True, i, 4.0 + 0.1 * i)
                r_g = bk.add_measurement(tc, pix[i], abi.FEAT, feat_R, True, i, 4.0 + 0.1 * i)
                assert r_o == r_g, f"step {k} feature {i}: add_measurement {r_g} vs oracle {r_o}"
                if not replace:
                    assert r_g == (abi.MEAS_NEW_FEATURE if k == 0 else abi.MEAS_SUCCESS)
        if check_each:
            _check(f, bk, f"step {k} before handle_measurements")
        gated_o = f.handle_measurements()
        ty, ids, res = bk.handle_measurements()
        if check_each:
            _check(f, bk, f"step {k} after handle_measurements ({len(ty)} updates)")
        gated_g = [int(ids[j]) for j in range(len(ty)) if ty[j] == abi.FEAT and res[j, 0] == abi.MEAS_GATED]
        assert gated_g == gated_o, f"step {k}: gated ids {gated_g} vs oracle {gated_o}"
        applied.append(len(ty))
        assert bk.get_time(0, 1)[0] == f.time()
    return f, bk, applied


def _check(f, bk, tag):
    ln, ids = bk.get_features()
    x, P = bk.get_state(), bk.get_covariance()
    for b in range(bk.B):
        assert ln[b] == f.get_len_features()
        assert list(ids[b][: ln[b]]) == f.tracked_features()
        ex, eP = tapes.live_compare(x[b], P[b], ln[b], f.get_state(), f.get_covariance(), f.get_len_features(), REL_TOL, f"{tag} inst {b}")
    fl, ctr = bk.get_flags()
    assert fl[0] == f.flags() and list(ctr[0]) == list(f.counters())
    return ex, eP


def test_queue_imu_acc_only(pkg, oracle_mod):
    """No camera: every ACC measurement is stamped at the newest input and is therefore fused one sample late."""
    f, bk, applied = _drive(pkg, oracle_mod, NF=0, n_imu=60, frame_every=10 ** 9, delay=0.0)
    assert sum(applied) >= 55
    _check(f, bk, "acc-only")


def test_queue_delayed_frames(pkg, oracle_mod):
    """Frames stamped 30 ms in the past: every frame rewinds the ring by several slots and replays inputs + ACC updates."""
    f, bk, applied = _drive(pkg, oracle_mod, NF=6, n_imu=120, frame_every=8, delay=0.0302)
    assert max(applied) >= 6
    _check(f, bk, "delayed frames")


def test_queue_late_measurement_and_short_ring(pkg, oracle_mod):
    """A ring of 12 slots wraps many times; ALT arrives older than already-handled measurements (not re-applied on replay)."""
    f, bk, applied = _drive(pkg, oracle_mod, NF=4, n_imu=150, frame_every=7, delay=0.0101, late_alt=True, hist=12)
    _check(f, bk, "late alt, short ring")


def test_queue_feature_drops(pkg, oracle_mod):
    """keep_only_features while frames are delayed: clear_feature compacts the current slot only (vi_ekf_feat.cpp:50-78), so the
    next rewind reads the older, uncompacted slot with the new feature count -- reproduced literally."""
    f, bk, applied = _drive(pkg, oracle_mod, NF=7, n_imu=160, frame_every=8, delay=0.0302, drop_every=3)
    assert f.get_len_features() < 7
    _check(f, bk, "drops")


@pytest.mark.parametrize("Qxf", [0.0, 1e-4])
def test_queue_new_features_behind_a_rewind(pkg, oracle_mod, Qxf):
    """A replacement feature arrives in a delayed frame: init_feature writes the current slot only (vi_ekf_feat.cpp:33-41), the
    next handle_measurements rewinds behind it, and the filter carries on with whatever the older slot held for that index --
    stale x, and for P the dead-slot block the reference keeps (P0_feat from init() plus Qx_feat per propagate, zeros after a
    clear_feature).  The engine maintains those dead rows in queue mode, so even this corner follows the reference."""
    f, bk, applied = _drive(pkg, oracle_mod, NF=5, n_imu=160, frame_every=8, delay=0.0302, drop_every=3, replace=True, Qxf=Qxf)
    assert max(f.tracked_features()) >= 5
    _check(f, bk, "new features behind a rewind")


def test_config0_vi_ekf_test_call_pattern(pkg, oracle_mod):
    """BASELINE configs[0]: the call sequence of the reference's test/vi_ekf_test.cpp:24-33,58 over the restated multirotor_sim
    stream (tests/queue_calls.py), N = 12 = the reference's release NUM_FEATURES: oracle vs engine, call for call."""
    from tests import queue_calls
    NF, B = 12, 2
    cfg = abi.default_config()
    calls = queue_calls.config0_calls(pkg, cfg, NF)
    f = oracle_mod.OracleVIEKF(cfg, NF)
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.enable_queue(250, 200)
    adds_o, gated_o, _ = queue_calls.replay(calls, f, abi, engine=False)
    adds_g, gated_g, n_upd = queue_calls.replay(calls, bk, abi, engine=True)
    assert adds_g == adds_o and gated_g == gated_o
    assert adds_g.count(abi.MEAS_NEW_FEATURE) >= NF and n_upd > 100 and f.get_len_features() == NF
    _check(f, bk, "configs[0] call pattern")


def test_queue_reset_and_errors(pkg):
    cfg = abi.default_config()
    bk = pkg.BatchVIEKF(cfg, B=2, N=2)
    with pytest.raises(pkg.batch.ViekfError):
        bk.add_measurement(0.0, [0.0], abi.ALT, [[1.0]])
    bk.enable_queue(8, 4)
    with pytest.raises(pkg.batch.ViekfError):
        bk.enable_queue(8, 4)
    u = np.array([0, 0, -9.8, 0, 0, 0.0])
    assert bk.add_measurement(0.5, [np.nan], abi.ALT, [[1.0]]) == abi.MEAS_NAN
    bk.queue_propagate_state(u, 1.0)
    assert bk.add_measurement(0.5, [2.0], abi.ALT, [[1.0]]) == abi.MEAS_INVALID
    bk.queue_propagate_state(u, 1.004)
    x1 = bk.get_state()
    bk.reset()
    assert bk.get_time(0, 1)[0] != 1.004
    bk.queue_propagate_state(u, 1.0)
    bk.queue_propagate_state(u, 1.004)
    assert np.array_equal(bk.get_state(), x1)


def test_queue_pending_measurement_of_a_dropped_feature(pkg, oracle_mod):
    """A FEAT measurement is queued, then keep_only_features drops its feature before handle_measurements replays it.  update()
    never initialises a feature (only add_measurement does, vi_ekf_meas.cpp:140-147): the replayed update is MEAS_INVALID with
    no state change and next_feature_id_ is not consumed, so later id assignments stay aligned with the oracle."""
    cfg = abi.default_config()
    cfg.use_keyframe_reset = 0
    NF, B = 4, 2
    f = oracle_mod.OracleVIEKF(cfg, NF)
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.enable_queue(250, 200)
    R = np.diag([0.25, 0.25])
    u = np.array([0.05, -0.02, -9.80665, 0.01, 0.0, -0.01])
    pix = {i: np.array([200.0 + 60 * i, 150.0 + 40 * i]) for i in range(3)}
    for k in range(12):
        t = 0.004 * (k + 1)
        f.propagate_state(u, t, True)
        bk.queue_propagate_state(u, t, True)
        if k == 0:
            for i in pix:
                assert f.add_measurement(t, pix[i], abi.FEAT, R, True, i, 5.0) == bk.add_measurement(t, pix[i], abi.FEAT, R, True, i, 5.0) == abi.MEAS_NEW_FEATURE
        if k == 6:
            for i in pix:  # stamped in the past: queued, not yet applied
                assert f.add_measurement(t - 0.006, pix[i] + 0.2, abi.FEAT, R, True, i) == bk.add_measurement(t - 0.006, pix[i] + 0.2, abi.FEAT, R, True, i) == abi.MEAS_SUCCESS
            f.keep_only_features([0, 2])
            bk.keep_only_features([0, 2])
            go = f.handle_measurements()
            ty, ids, res = bk.handle_measurements()
            assert list(ids) == [2, 1, 0] and list(res[:, 0]) == [abi.MEAS_SUCCESS, abi.MEAS_INVALID, abi.MEAS_SUCCESS] and go == []
        if k == 8:  # the next new feature gets the same internal id on both sides
            z = np.array([400.0, 300.0])
            assert f.add_measurement(t, z, abi.FEAT, R, True, 7, 6.0) == bk.add_measurement(t, z, abi.FEAT, R, True, 7, 6.0) == abi.MEAS_NEW_FEATURE
    _check(f, bk, "pending measurement of a dropped feature")
    assert f.tracked_features() == [0, 2, 3]
