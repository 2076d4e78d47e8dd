"""The N = 0 executor (csrc/viekf_octet.cu: eight lanes per filter, four filters per warp, covariance in registers) against the CPU
oracle on the paths the shared tests do not reach at N = 0: per-instance values that make the filters of one warp take different
gate decisions, a NaN measurement on some instances only, ragged batch sizes (lanes and warps without a filter), every base
measurement model under every configuration variant, keyframe bookkeeping through KEEP with nothing tracked, chunked launches,
sharding, and the warp-per-filter mapping of the shared-memory executor as a second opinion.

Bit-exact: per-event meas_result_t, flags, counters.  x / P within 1e-9 relative (REL_TOL, the north-star tolerance)."""
import numpy as np
import pytest

from tests.common import REL_TOL, abi
from tests import tapes
from tests.test_gpu_parity import _compare, _default_cfg, _stats_close

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("drag,partial,kfr", [(1, 1, 1), (0, 0, 0), (1, 0, 1), (0, 1, 0)])
@pytest.mark.parametrize("B", [1, 5, 37])
def test_all_base_models_and_ragged_batches(pkg, oracle_mod, drag, partial, kfr, B):
    cfg = _default_cfg(use_drag_term=drag, use_partial_update=partial, use_keyframe_reset=kfr)
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, 0, steps=70, seed=11 + B)
    ev2, vals = tapes.perturbed_values(ev, B, seed=5, scale=2.0)
    ref = oracle_mod.run_batch(cfg, 0, ev2, idp, values=vals, count=B, nthreads=8, n_buckets=nb)
    bk = pkg.BatchVIEKF(cfg, B=B, N=0)
    bk.stats_reset(nb)
    res = bk.run_tape(ev2, idp, values=vals, results_first=0, results_count=B)
    assert np.array_equal(res, ref["results"])
    kinds = set(int(r) for r in np.unique(res))
    assert abi.MEAS_SUCCESS in kinds and (abi.MEAS_GATED in kinds or not drag)
    _compare(bk, ref, tag=f"octet drag={drag} partial={partial} kfr={kfr} B={B}")
    _stats_close(bk.stats(nb), ref["stats"])
    P = bk.get_covariance()
    assert np.array_equal(P, np.transpose(P, (0, 2, 1))), "the register rows must stay symmetric bit for bit"


def test_filters_of_one_warp_diverge(pkg, oracle_mod):
    """Neighbouring instances (the four filters of a warp) gate differently, and some receive a NaN measurement."""
    cfg = _default_cfg()
    B = 16
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, 0, steps=120, seed=29, outlier_every=0)
    ev2, vals = tapes.perturbed_values(ev, B, seed=6, scale=1.0)
    rng = np.random.default_rng(0)
    acc = [i for i in range(len(ev2)) if ev2[i]["kind"] == abi.EV_UPD and ev2[i]["type"] == abi.ACC]
    for i in acc[::3]:  # outliers on a random subset of instances: gate decisions differ inside a warp
        nbk = int(ev2[i]["noise"])
        hit = rng.uniform(size=B) < 0.3
        vals[hit, 8 * nbk] += 40.0
    for i in acc[5::7]:  # NaN in z on two instances of the same warp: rejected by add_measurement (MEAS_NAN)
        nbk = int(ev2[i]["noise"])
        vals[[1, 2], 8 * nbk + 1] = np.nan
    ref = oracle_mod.run_batch(cfg, 0, ev2, idp, values=vals, count=B, nthreads=8, n_buckets=nb)
    bk = pkg.BatchVIEKF(cfg, B=B, N=0)
    bk.stats_reset(nb)
    res = bk.run_tape(ev2, idp, values=vals, results_first=0, results_count=B)
    assert np.array_equal(res, ref["results"])
    assert (res == abi.MEAS_NAN).sum() > 0 and (res == abi.MEAS_GATED).sum() > 0
    quad = res[0:4]
    assert len({tuple(r) for r in quad}) == 4, "the four filters of the first warp should all take different decisions"
    _compare(bk, ref, tag="octet divergence")
    _stats_close(bk.stats(nb), ref["stats"])


def test_feature_events_with_nothing_tracked(pkg, oracle_mod):
    """FEAT / DEPTH / INIT / CLEAR / KEEP events on a filter that can track nothing: results as the oracle's, state untouched, the
    keyframe list and the keyframe reset still follow keep_only_features (vi_ekf_feat.cpp:81-142)."""
    cfg = _default_cfg(use_keyframe_reset=1)
    B = 6
    tb = pkg.TapeBuilder()
    rng = np.random.default_rng(4)
    t = 0.0
    tb.propagate(tapes.imu_at(t), t)
    for k in range(1, 41):
        t = 0.005 * k
        tb.propagate(tapes.imu_at(t, rng, 0.05), t)
        if k % 6 == 0:
            tb.keep_only_features([3, 4, 100 + k])
            tb.update(abi.FEAT, np.array([300.0, 200.0]), np.eye(2) * 0.25, id=k, depth=4.0)
            tb.update(abi.DEPTH, np.array([3.0]), np.array([[0.5]]), id=k)
            tb.init_feature(np.array([310.0, 210.0]), 900 + k, 7.0)
            tb.clear_feature(k)
        if k % 10 == 0:
            tb.keyframe_reset()
        if k % 8 == 0:
            tb.update(abi.ALT, np.array([0.1 + 0.01 * k]), np.array([[0.1]]))
    ev, idp = tb.events(), tb.id_pool()
    ref = oracle_mod.run_batch(cfg, 0, ev, idp, count=B, nthreads=4)
    bk = pkg.BatchVIEKF(cfg, B=B, N=0)
    res = bk.run_tape(ev, idp, results_first=0, results_count=B)
    assert np.array_equal(res, ref["results"])
    assert (res == abi.MEAS_NEW_FEATURE).sum() > 0 and (res == abi.MEAS_INVALID).sum() > 0
    _compare(bk, ref, tag="octet feature events")
    pose, gcov = bk.get_global_pose()
    assert np.abs(pose[0, :3]).max() > 0, "keep_only_features / keyframe_reset should have moved the node"
    ln, ids = bk.get_features()
    assert (ln == 0).all()


def test_chunked_sharded_and_other_mapping(pkg, oracle_mod, monkeypatch):
    cfg = _default_cfg()
    B = 22
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, 0, steps=64, seed=17)
    for i in range(len(ev)):
        if ev[i]["kind"] == abi.EV_PROP:
            ev[i]["sigma"][:6] = [0.05, 0.05, 0.05, 0.002, 0.002, 0.002]
    bias = [0.02, 0.02, 0.02, 0.002, 0.002, 0.002]
    seed = 0xC0FFEE
    a = pkg.BatchVIEKF(cfg, B=B, N=0)
    a.perturb_x0(5)
    ra = a.run_tape(ev, idp, results_first=0, results_count=B, seed=seed, bias_range=bias)
    xa, Pa = a.get_state(), a.get_covariance()
    # chunked = single launch, bit for bit
    b = pkg.BatchVIEKF(cfg, B=B, N=0)
    b.perturb_x0(5)
    d = b.upload_tape(ev, idp, seed=seed, bias_range=bias)
    cuts = [0, 1, 9, 10, len(ev) // 2, len(ev)]
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        b.run_dev_tape(d, lo, hi - lo, sync=False)
    b.sync()
    b.free_tape(d)
    assert np.array_equal(b.get_state(), xa) and np.array_equal(b.get_covariance(), Pa)
    # sharded = unsharded, bit for bit (instance offsets that do not fall on a warp boundary)
    parts = []
    for off, cnt in ((0, 7), (7, 15)):
        s = pkg.BatchVIEKF(cfg, B=cnt, N=0)
        s.set_instance_offset(off)
        s.perturb_x0(5)
        s.run_tape(ev, idp, seed=seed, bias_range=bias)
        parts.append((s.get_state(), s.get_covariance()))
    assert np.array_equal(np.concatenate([p[0] for p in parts]), xa)
    assert np.array_equal(np.concatenate([p[1] for p in parts]), Pa)
    # the warp-per-filter mapping of the shared-memory executor: same decisions, numbers within the tolerance of each other
    monkeypatch.setenv("VIEKF_GROUP", "32")
    c = pkg.BatchVIEKF(cfg, B=B, N=0)
    c.perturb_x0(5)
    rc = c.run_tape(ev, idp, results_first=0, results_count=B, seed=seed, bias_range=bias)
    assert np.array_equal(ra, rc)
    xc, Pc = c.get_state(), c.get_covariance()
    for i in range(B):
        tapes.live_compare(xa[i], Pa[i], 0, xc[i], Pc[i], 0, REL_TOL, f"octet vs warp-per-filter inst {i}")
