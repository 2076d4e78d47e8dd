"""Parity of the CUDA engine (through the C ABI) against the CPU oracle on the same seeded tapes.

Tolerance (BASELINE.json north_star): feature indexing, keep/drop and gate decisions bit-exact; state and covariance
within a relative 1e-9 (state: |dx| / max(1,|x|); covariance: |dP_ij| / sqrt(P_ii P_jj)).
"""
import numpy as np
import pytest

from tests.common import REL_TOL, abi, jac_test_config
from tests import tapes

pytestmark = pytest.mark.gpu


def _default_cfg(**kw):
    cfg = abi.default_config()
    for k, v in kw.items():
        setattr(cfg, k, v)
    return cfg


def _stats_close(a, b, rtol=1e-7):
    """Statistics blocks agree: same NaN pattern (NEES is undefined right after a keyframe reset), values within rtol."""
    assert a.shape == b.shape
    assert np.array_equal(np.isnan(a), np.isnan(b)), f"NaN pattern differs at {np.argwhere(np.isnan(a) != np.isnan(b))[:5].tolist()}"
    d = np.abs(np.nan_to_num(a) - np.nan_to_num(b)) / np.maximum(np.abs(np.nan_to_num(b)), 1e-2)
    d[:, 33:35] *= 1e-3  # NEES = e^T P^-1 e amplifies the 1e-9 agreement of P by its condition number (P0 spans 1e-6 .. 0.3)
    i = np.unravel_index(np.argmax(d), d.shape)
    assert d[i] <= rtol, f"stats differ at {i}: {a[i]!r} vs {b[i]!r}"


def _compare(bk, ref, idxs=None, tol=REL_TOL, tag=""):
    """bk: BatchVIEKF after a run; ref: oracle.run_batch dict."""
    B = bk.B
    idxs = range(B) if idxs is None else idxs
    x = bk.get_state()
    P = bk.get_covariance()
    ln, ids = bk.get_features()
    fl, ctr = bk.get_flags()
    worst = (0.0, 0.0)
    for b in idxs:
        assert ln[b] == ref["len"][b], f"{tag} inst {b}: len {ln[b]} vs {ref['len'][b]}"
        assert list(ids[b][: ln[b]]) == list(ref["ids"][b][: ln[b]]), f"{tag} inst {b}: feature ids differ"
        assert fl[b] == ref["flags"][b, 0], f"{tag} inst {b}: flags {fl[b]} vs {ref['flags'][b, 0]}"
        assert list(ctr[b]) == list(ref["flags"][b, 1:4]), f"{tag} inst {b}: counters {ctr[b]} vs {ref['flags'][b, 1:4]}"
        ex, eP = tapes.live_compare(x[b], P[b], ln[b], ref["x"][b], ref["P"][b], ref["len"][b], tol, f"{tag} inst {b}")
        worst = (max(worst[0], ex), max(worst[1], eP))
    return worst


@pytest.mark.parametrize("NF,steps", [(0, 40), (3, 60), (12, 72), (35, 48)])
def test_tape_parity_all_event_kinds(pkg, oracle_mod, NF, steps):
    cfg = _default_cfg()
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=steps, seed=10 + NF)
    B = 5
    ref = oracle_mod.run_batch(cfg, NF, ev, idp, count=B, nthreads=4, n_buckets=nb)
    # the tape-writing filter and the batch runner are the same oracle: sanity
    assert np.array_equal(ref["x"][0], f.get_state())
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.stats_reset(nb)
    res = bk.run_tape(ev, idp, results_first=0, results_count=B)
    assert np.array_equal(res, ref["results"]), "meas_result_t per event differ"
    kinds = set(int(r) for r in np.unique(res))
    if NF > 0:
        assert {abi.MEAS_SUCCESS, abi.MEAS_GATED, abi.MEAS_NEW_FEATURE}.issubset(kinds)
    _compare(bk, ref, tag=f"N={NF}")
    st = bk.stats(nb)
    assert np.allclose(st[:, 0], B)
    _stats_close(st, ref["stats"])
    if NF > 0:
        pose, gcov = bk.get_global_pose()
        t7, c36 = f.global_pose()
        # node pose of the keyframe resets (vi_ekf_kfr.cpp:136-140): compare through the oracle's global pose of the nominal instance
        x = bk.get_state()[0]
        nt, nq = pose[0, :3], pose[0, 3:]
        gt = nt + oracle_mod.quat_op(0, nq, x[0:3], 3)
        assert np.allclose(gt, t7[:3], rtol=1e-9, atol=1e-9)


@pytest.mark.parametrize("drag,partial,kfr", [(0, 0, 0), (1, 0, 1), (0, 1, 0)])
def test_tape_parity_config_variants(pkg, oracle_mod, drag, partial, kfr):
    cfg = _default_cfg(use_drag_term=drag, use_partial_update=partial, use_keyframe_reset=kfr)
    NF = 5
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=48, seed=77 + drag + 2 * partial)
    ref = oracle_mod.run_batch(cfg, NF, ev, idp, count=2, nthreads=2)
    bk = pkg.BatchVIEKF(cfg, B=2, N=NF)
    res = bk.run_tape(ev, idp, results_first=0, results_count=2)
    assert np.array_equal(res, ref["results"])
    _compare(bk, ref, tag=f"drag={drag} partial={partial} kfr={kfr}")


def test_random_jac_test_config(pkg, oracle_mod):
    """The reference's init_jacobians_test recipe (test/jac_test.cpp:118-170, non-unit q_b_c) as filter configuration."""
    rng = np.random.default_rng(5)
    cfg, x0, att = jac_test_config(rng)
    x0[6:10] = oracle_mod.quat_op(6, x0[6:10], att * 0.2)
    cfg.x0[:] = list(x0)
    NF = 6
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=30, seed=3, extra_models=False, outlier_every=0)
    ref = oracle_mod.run_batch(cfg, NF, ev, idp, count=1)
    bk = pkg.BatchVIEKF(cfg, B=1, N=NF)
    res = bk.run_tape(ev, idp, results_first=0, results_count=1)
    assert np.array_equal(res, ref["results"])
    # unit P0 collapsing to ~1e-6 under 0.5 px measurements: each covariance-form update loses eps*|P_old|/|P_new| relative
    # accuracy whatever the summation order, so two correct evaluations agree to a few 1e-9 here, not 1e-9
    _compare(bk, ref, tol=5e-9, tag="jac_test config")


def test_per_instance_values(pkg, oracle_mod):
    cfg = _default_cfg()
    NF, B = 8, 37  # ragged against the group / CTA size
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=60, seed=21)
    ev2, vals = tapes.perturbed_values(ev, B, seed=4)
    x0 = np.tile(np.array(cfg.x0[:]), (B, 1))
    x0[:, 0:3] += np.random.default_rng(2).normal(size=(B, 3)) * 0.01
    ref = oracle_mod.run_batch(cfg, NF, ev2, idp, values=vals, x0=x0, count=B, nthreads=8, n_buckets=nb)
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.set_x0(x0)
    bk.stats_reset(nb)
    res = bk.run_tape(ev2, idp, values=vals, results_first=0, results_count=B)
    assert np.array_equal(res, ref["results"])
    assert len({tuple(r) for r in res}) > 1, "perturbed instances should not all take the same decisions"
    _compare(bk, ref, tag="values")
    _stats_close(bk.stats(nb), ref["stats"])


def test_generated_noise_and_sharding(pkg, oracle_mod):
    """Philox noise drawn on the device equals the oracle's host draw; a batch split over two handles with instance
    offsets reproduces the unsplit batch bit for bit (the multi-GPU sharding rule)."""
    cfg = _default_cfg()
    NF, B = 4, 24
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=48, seed=8, extra_models=False)
    for i in range(len(ev)):
        if ev[i]["kind"] == abi.EV_PROP:
            ev[i]["sigma"][:6] = [0.05, 0.05, 0.05, 0.002, 0.002, 0.002]
        elif ev[i]["kind"] == abi.EV_UPD and ev[i]["type"] == abi.FEAT:
            ev[i]["sigma"][:2] = 0.3
    bias = [0.02, 0.02, 0.02, 0.002, 0.002, 0.002]
    seed = 0x5EED
    ref = oracle_mod.run_batch(cfg, NF, ev, idp, count=B, nthreads=8, seed=seed, bias_range=bias, perturb_seed=99, n_buckets=nb)
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.perturb_x0(99)
    bk.stats_reset(nb)
    res = bk.run_tape(ev, idp, results_first=0, results_count=B, seed=seed, bias_range=bias)
    assert np.array_equal(res, ref["results"])
    _compare(bk, ref, tag="noise")
    _stats_close(bk.stats(nb), ref["stats"])
    xa, Pa = bk.get_state(), bk.get_covariance()
    h = B // 2
    parts = []
    for off in (0, h):
        s = pkg.BatchVIEKF(cfg, B=h, N=NF)
        s.set_instance_offset(off)
        s.perturb_x0(99)
        s.run_tape(ev, idp, seed=seed, bias_range=bias)
        parts.append((s.get_state(), s.get_covariance()))
    assert np.array_equal(np.concatenate([parts[0][0], parts[1][0]]), xa)
    assert np.array_equal(np.concatenate([parts[0][1], parts[1][1]]), Pa)


def test_per_call_api_matches_oracle(pkg, oracle_mod):
    """The reference-shaped calls, one launch each: propagate_state / update / init_feature / keep_only_features /
    clear_feature / keyframe_reset (reference include/vi_ekf.h:294-309)."""
    cfg = _default_cfg()
    NF, B = 4, 3
    rng = np.random.default_rng(12)
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    fs = [oracle_mod.OracleVIEKF(cfg, NF, hist=2) for _ in range(B)]
    t = 0.0
    u = tapes.imu_at(t)
    bk.propagate_state(u, t)
    for f in fs:
        f.propagate_state(u, t, False)
    for i in range(NF + 1):  # one more than fits: refused
        pix = rng.uniform(100, 400, size=(B, 2))
        ok = bk.init_feature(pix, i, depth=5.0 + i)
        exp = [fs[b].init_feature(pix[b], i, 5.0 + i) for b in range(B)]
        assert list(ok) == exp
    for k in range(1, 25):
        t = 0.004 * k
        ub = np.stack([tapes.imu_at(t, rng, 0.05) for _ in range(B)])
        bk.propagate_state(ub, t)
        for b in range(B):
            fs[b].propagate_state(ub[b], t, False)
        z = np.stack([fs[b].h_model(abi.ACC, fs[b].get_state())[0][:2] + rng.normal(size=2) * 0.3 for b in range(B)])
        r = bk.update(abi.ACC, z, np.eye(2))
        assert list(r) == [fs[b].update_direct(abi.ACC, z[b], np.eye(2)) for b in range(B)]
        if k % 5 == 0:
            for fid in reversed(fs[0].tracked_features()):
                z = np.stack([fs[b].h_model(abi.FEAT, fs[b].get_state(), fid)[0][:2] + rng.normal(size=2) * 0.4 for b in range(B)])
                r = bk.update(abi.FEAT, z, np.eye(2) * 0.25, id=fid)
                assert list(r) == [fs[b].update_direct(abi.FEAT, z[b], np.eye(2) * 0.25, id=fid) for b in range(B)]
        if k == 12:
            keep = fs[0].tracked_features()[1:]
            bk.keep_only_features(keep)
            for f in fs:
                f.keep_only_features(keep)
        if k == 14:
            z = rng.uniform(100, 400, size=(B, 2))
            r = bk.update(abi.FEAT, z, np.eye(2) * 0.25, id=fs[0].tracked_features()[-1] + 1)
            assert list(r) == [abi.MEAS_NEW_FEATURE] * B
            for b in range(B):
                fs[b].update_direct(abi.FEAT, z[b], np.eye(2) * 0.25, id=fs[b].tracked_features()[-1] + 1)
        if k == 18:
            bk.clear_feature(fs[0].tracked_features()[0])
            for f in fs:
                f.clear_feature(f.tracked_features()[0])
        if k == 20:
            bk.keyframe_reset()
            for f in fs:
                f.keyframe_reset()
    x, P = bk.get_state(), bk.get_covariance()
    ln, ids = bk.get_features()
    tt = bk.get_time()
    for b in range(B):
        assert list(ids[b][: ln[b]]) == fs[b].tracked_features()
        assert tt[b] == fs[b].time()
        tapes.live_compare(x[b], P[b], ln[b], fs[b].get_state(), fs[b].get_covariance(), fs[b].get_len_features(), REL_TOL, f"per-call inst {b}")
    assert bk.launch_count() > 40


def test_chunked_tape_equals_single_launch(pkg, oracle_mod):
    cfg = _default_cfg()
    NF, B = 6, 9
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=40, seed=31)
    a = pkg.BatchVIEKF(cfg, B=B, N=NF)
    a.run_tape(ev, idp)
    b = pkg.BatchVIEKF(cfg, B=B, N=NF)
    d = b.upload_tape(ev, idp)
    cuts = [0, 7, 8, 50, len(ev) // 2, len(ev)]
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        b.run_dev_tape(d, lo, hi - lo, sync=False)
    b.sync()
    b.free_tape(d)
    assert np.array_equal(a.get_state(), b.get_state())
    assert np.array_equal(a.get_covariance(), b.get_covariance())
    assert b.launch_count() == len(cuts) - 1


@pytest.mark.parametrize("group", [1, 2, 32, 64, 128, 256])
def test_every_thread_group_mapping(pkg, oracle_mod, group, monkeypatch):
    """The same filter under each mapping (1 = register-tile executor, 2 = fragment executor, warp / 2 warps / CTA of the shared-memory
    executor) stays within tolerance."""
    monkeypatch.setenv("VIEKF_GROUP", str(group))
    cfg = _default_cfg()
    NF = 7
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=36, seed=41)
    ref = oracle_mod.run_batch(cfg, NF, ev, idp, count=6, nthreads=6)
    bk = pkg.BatchVIEKF(cfg, B=6, N=NF)
    res = bk.run_tape(ev, idp, results_first=0, results_count=6)
    assert np.array_equal(res, ref["results"])
    _compare(bk, ref, tag=f"group={group}")


@pytest.mark.parametrize("fpc", [1, 2, 6])
def test_tile_executor_filters_per_cta(pkg, oracle_mod, fpc, monkeypatch):
    """The register-tile executor with 1 / 2 / 6 filters per CTA in event lock-step (B = 6 leaves groups without a filter in
    the last CTA) gives the same decisions and, bit for bit, the same numbers."""
    monkeypatch.setenv("VIEKF_TILE_FPC", str(fpc))
    cfg = _default_cfg()
    NF, B = 9, 6
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=48, seed=13)
    ref = oracle_mod.run_batch(cfg, NF, ev, idp, count=B, nthreads=6, n_buckets=nb)
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.stats_reset(nb)
    res = bk.run_tape(ev, idp, results_first=0, results_count=B)
    assert np.array_equal(res, ref["results"])
    _compare(bk, ref, tag=f"fpc={fpc}")
    _stats_close(bk.stats(nb), ref["stats"])
    x, P = bk.get_state(), bk.get_covariance()
    monkeypatch.setenv("VIEKF_TILE_FPC", "1")
    b1 = pkg.BatchVIEKF(cfg, B=B, N=NF)
    b1.run_tape(ev, idp)
    assert np.array_equal(b1.get_state(), x) and np.array_equal(b1.get_covariance(), P)


def test_edge_cases(pkg, oracle_mod):
    cfg = _default_cfg()
    bk = pkg.BatchVIEKF(cfg, B=2, N=2)
    # empty tape is a no-op
    tb = pkg.TapeBuilder()
    bk.run_tape(tb.events(), tb.id_pool())
    x0 = bk.get_state()
    assert np.allclose(x0[0, :17], np.array(cfg.x0[:]))
    # NaN measurement -> MEAS_NAN, nothing changes (vi_ekf_meas.cpp:136-137)
    bk.propagate_state(tapes.imu_at(0.0), 0.0)
    r = bk.update(abi.ALT, np.array([float("nan")]), np.array([[0.1]]))
    assert list(r) == [abi.MEAS_NAN] * 2
    # backwards propagation is refused and flagged (vi_ekf.cpp:285-289); |dt| < 1e-6 is a no-op (:282)
    bk.propagate_state(tapes.imu_at(0.01), 0.01)
    xa = bk.get_state()
    bk.propagate_state(tapes.imu_at(0.01), 0.01 + 5e-7)
    assert np.array_equal(bk.get_state(), xa)
    bk.propagate_state(tapes.imu_at(0.0), 0.005)
    assert np.array_equal(bk.get_state(), xa)
    fl, _ = bk.get_flags()
    assert all(f & abi.FLAG_BACKWARD_DT for f in fl)
    # negative inverse depth is repaired by fix_depth (vi_ekf_helper.cpp:142-148)
    assert all(bk.init_feature(np.array([300.0, 200.0]), 0, depth=4.0))
    r = bk.update(abi.INV_DEPTH, np.array([-50.0]), np.array([[1e-6]]), id=0, active=True)
    f = oracle_mod.OracleVIEKF(cfg, 2, hist=2)
    f.propagate_state(tapes.imu_at(0.0), 0.0, False)
    f.propagate_state(tapes.imu_at(0.01), 0.01, False)
    f.init_feature(np.array([300.0, 200.0]), 0, 4.0)
    ro = f.update_direct(abi.INV_DEPTH, np.array([-50.0]), np.array([[1e-6]]), id=0)
    assert list(r) == [ro] * 2
    fl, ctr = bk.get_flags()
    assert (int(fl[0]) & ~abi.FLAG_BACKWARD_DT) == f.flags()
    x, P = bk.get_state(), bk.get_covariance()
    tapes.live_compare(x[0], P[0], 1, f.get_state(), f.get_covariance(), 1, REL_TOL, "fix_depth")
    # argument errors are reported, not crashed on
    with pytest.raises(pkg.batch.ViekfError):
        bk.get_state(first=1, count=5)
    with pytest.raises(pkg.batch.ViekfError):
        pkg.BatchVIEKF(cfg, B=0, N=2)


def test_thousand_steps_25_features(pkg, oracle_mod):
    """north_star: state and covariance within 1e-9 relative over 1000 steps, decisions bit-exact (N = 25)."""
    cfg = _default_cfg()
    NF, B = 25, 8
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=1000, seed=1234, cam_every=7, churn_every=1, extra_models=False,
                                           outlier_every=23, stat_every=100)
    for i in range(len(ev)):
        if ev[i]["kind"] == abi.EV_PROP:
            ev[i]["sigma"][:6] = [0.02, 0.02, 0.02, 0.001, 0.001, 0.001]
        elif ev[i]["kind"] == abi.EV_UPD and ev[i]["type"] == abi.FEAT:
            ev[i]["sigma"][:2] = 0.2
    seed = 4242
    ref = oracle_mod.run_batch(cfg, NF, ev, idp, count=B, nthreads=8, seed=seed, n_buckets=nb)
    # decisions must not sit on a threshold for the comparison to be meaningful (hard parts)
    assert ref["margins"][:, 0].min() > 1e-7, "a gate decision sits within 1e-7 of the threshold: pick another seed"
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.stats_reset(nb)
    res = bk.run_tape(ev, idp, results_first=0, results_count=B, seed=seed)
    assert np.array_equal(res, ref["results"])
    worst = _compare(bk, ref, tag="1000 steps N=25")
    print(f"1000 steps, N=25: worst state rel err {worst[0]:.2e}, covariance rel err {worst[1]:.2e}")
    _stats_close(bk.stats(nb), ref["stats"], rtol=1e-6)


@pytest.mark.parametrize("B,NF,steps", [(65536, 0, 50), (65536, 8, 14), (262144, 25, 8)])
def test_baseline_sizes_properties(pkg, oracle_mod, B, NF, steps):
    """BASELINE.json configs #2-#4 at full batch size: size-independent properties plus an oracle check of a sample."""
    cfg = _default_cfg()
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=steps, seed=5, cam_every=7, churn_every=1, extra_models=False, stat_every=0)
    for i in range(len(ev)):
        if ev[i]["kind"] == abi.EV_PROP:
            ev[i]["sigma"][:6] = [0.02, 0.02, 0.02, 0.001, 0.001, 0.001]
    seed = 17
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.run_tape(ev, idp, seed=seed)
    sample = [0, 1, B // 3, B - 2, B - 1]
    ln, ids = bk.get_features()
    fl, ctr = bk.get_flags()
    assert (fl & abi.FLAG_NAN_STATE).sum() == 0
    assert (ln == ln[0]).all(), "the tape is shared: every instance tracks the same number of features"
    assert (ids == ids[0]).all()
    # idempotence of the read-out and determinism of the launch
    bk2 = pkg.BatchVIEKF(cfg, B=4096, N=NF)
    bk2.run_tape(ev, idp, seed=seed)
    assert np.array_equal(bk2.get_state(), bk.get_state(0, 4096))
    for b in sample:
        ref = oracle_mod.run_batch(cfg, NF, ev, idp, count=1, seed=seed, inst_off=b)
        x = bk.get_state(b, 1)[0]
        P = bk.get_covariance(b, 1)[0]
        assert np.allclose(P, P.T)
        d = np.diag(P)[: 16 + 3 * ln[b]]
        assert (d >= 0).all() and (d[3:] > 0).all()  # the position block is exactly zero right after a keyframe reset
        tapes.live_compare(x, P, ln[b], ref["x"][0], ref["P"][0], ref["len"][0], REL_TOL, f"B={B} inst {b}")


def test_read_back_getters(pkg, oracle_mod):
    """get_depths / get_zetas / get_qzetas / get_zeta / get_feat / get_depth / get_covariance_diagonal, set_imu_bias, set_drag_term
    (reference include/vi_ekf.h:276-292, src/vi_ekf/vi_ekf.cpp:180-260) against the oracle's state after a tape."""
    from oracle import np_oracle
    cfg = _default_cfg()
    NF, B = 6, 3
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=30, seed=9, extra_models=False)
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.run_tape(ev, idp)
    x, P = f.get_state(), f.get_covariance()
    ln = f.get_len_features()
    ids = f.tracked_features()
    assert 0 < ln <= NF
    dep, zet, qz = bk.get_depths(), bk.get_zetas(), bk.get_qzetas()
    dg = bk.get_covariance_diagonal()
    for b in range(B):
        for i in range(NF):
            if i < ln:
                q = x[17 + 5 * i:17 + 5 * i + 4]
                assert np.allclose(dep[b, i], 1.0 / x[17 + 5 * i + 4], rtol=1e-9)
                assert np.allclose(qz[b, i], q, rtol=1e-9, atol=1e-12)
                assert np.allclose(zet[b, i], np_oracle.zeta_of(q), rtol=1e-9, atol=1e-12)
            else:
                assert np.isnan(dep[b, i]) and np.isnan(zet[b, i]).all() and np.isnan(qz[b, i]).all()
        n = 16 + 3 * ln
        assert np.allclose(dg[b, :n], np.diag(P)[:n], rtol=1e-9, atol=1e-15) and (dg[b, n:] == 0).all()
    assert np.array_equal(bk.get_zeta(0), zet[:, 0])
    fid = ids[-1]
    zh, _ = f.h_model(abi.FEAT, x, fid)
    assert np.allclose(bk.get_feat(fid), np.tile(zh[:2], (B, 1)), rtol=1e-9)
    assert np.allclose(bk.get_depth(fid), 1.0 / x[17 + 5 * (ln - 1) + 4], rtol=1e-9)
    assert np.isnan(bk.get_depth(123456)).all() and np.isnan(bk.get_feat(123456)).all()
    # set_imu_bias / set_drag_term
    bk.set_imu_bias([0.01, 0.02, 0.03], [0.1, 0.2, 0.3])
    xs = bk.get_state()
    assert np.array_equal(xs[:, 13:16], np.tile([0.01, 0.02, 0.03], (B, 1))) and np.array_equal(xs[:, 10:13], np.tile([0.1, 0.2, 0.3], (B, 1)))
    assert bk.get_drag_term() is True
    bk.set_drag_term(False)
    assert bk.get_drag_term() is False
    cfg2 = _default_cfg(use_drag_term=0)
    # a no-drag propagate from the same state: the handle now follows the no-drag dynamics (vi_ekf_dyn.cpp:44-49)
    t0 = bk.get_time(0, 1)[0]
    u = tapes.imu_at(t0 + 0.005)
    bk.propagate_state(u, t0 + 0.005)
    x1 = bk.get_state()[0]
    g2 = np_oracle.NpVIEKF(cfg2, NF)
    g2.x, g2.P = xs[0].copy(), bk.get_covariance()[0].copy()
    g2.len, g2.ids, g2.started, g2.t = ln, list(ids), True, t0
    g2.propagate_state(u, t0 + 0.005)
    assert np.allclose(x1[:17 + 5 * ln], g2.x[:17 + 5 * ln], rtol=1e-9, atol=1e-12)
