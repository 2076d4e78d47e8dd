"""The fragment executor (csrc/viekf_frag.cu: one warp per filter, covariance in FP64 tensor-core accumulator fragments, 1 <= N <= 8)
against the CPU oracle where the shared tests leave gaps: every capacity N = 1 ... 8 (the live-block guards below N = 6, the unguarded
instantiation from 6 on), batches that do not fill a CTA of eight filters, the three fix_depth repairs, a refused init_feature, feature
removal in the middle, chunked / sharded bit-identity at N = 8, and the round-1 mappings as a second opinion.

Bit-exact: per-event meas_result_t, ids, flags, counters.  x / P within 1e-9 relative (REL_TOL, the north-star tolerance)."""
import numpy as np
import pytest

from tests.common import REL_TOL, abi
from tests import tapes
from tests.test_gpu_parity import _compare, _default_cfg, _stats_close

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("NF", [1, 2, 4, 5, 6, 7, 8])
def test_every_capacity(pkg, oracle_mod, NF):
    cfg = _default_cfg()
    B = 11  # one full CTA of eight filters and a ragged one
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=72, seed=50 + NF, cam_every=5, churn_every=2)
    ev2, vals = tapes.perturbed_values(ev, B, seed=9)
    ref = oracle_mod.run_batch(cfg, NF, ev2, idp, values=vals, count=B, nthreads=8, n_buckets=nb)
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    bk.stats_reset(nb)
    res = bk.run_tape(ev2, idp, values=vals, results_first=0, results_count=B)
    assert np.array_equal(res, ref["results"])
    _compare(bk, ref, tag=f"frag N={NF}")
    _stats_close(bk.stats(nb), ref["stats"])
    P = bk.get_covariance()
    assert np.array_equal(P, np.transpose(P, (0, 2, 1)))


def test_fix_depth_repairs_refused_init_and_removal(pkg, oracle_mod):
    cfg = _default_cfg()
    NF, B = 3, 9
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
    f = oracle_mod.OracleVIEKF(cfg, NF, hist=2)

    def both(fn_engine, fn_oracle):
        re, ro = fn_engine(), fn_oracle()
        return re, ro

    def check(tag):
        x, P = bk.get_state(), bk.get_covariance()
        ln, ids = bk.get_features()
        assert ln[0] == f.get_len_features() and list(ids[0][: ln[0]]) == list(f.tracked_features())
        for b in (0, B - 1):
            tapes.live_compare(x[b], P[b], ln[b], f.get_state(), f.get_covariance(), f.get_len_features(), REL_TOL, tag)
        fl, ctr = bk.get_flags()
        assert int(fl[0]) == f.flags(), tag

    t = 0.0
    for k in range(3):
        t = 0.005 * k
        bk.propagate_state(tapes.imu_at(t), t)
        f.propagate_state(tapes.imu_at(t), t, False)
    # feature 1 starts at rho = 1 / 0.004 = 250 (> 100: reset + P(rho, rho) = P0_feat, vi_ekf_helper.cpp:149-154), feature 2 at
    # rho = 1 / -2 (< 0: reset + variance inflation, :142-148); the next propagate_state runs fix_depth
    depths = [3.0, 0.004, -2.0]
    for i in range(NF):
        pix = np.array([200.0 + 60 * i, 180.0 + 30 * i])
        assert all(bk.init_feature(pix, i, depth=depths[i]))
        f.init_feature(pix, i, depths[i])
    assert not any(bk.init_feature(np.array([100.0, 100.0]), 99, depth=2.0)), "a full filter refuses init_feature (vi_ekf_feat.cpp:9)"
    check("after init")
    t = 0.015
    bk.propagate_state(tapes.imu_at(t), t)
    f.propagate_state(tapes.imu_at(t), t, False)
    check("after fix_depth")
    fl, ctr = bk.get_flags()
    assert (fl[0] & abi.FLAG_DEPTH_BIG_FIXED) and (fl[0] & abi.FLAG_DEPTH_NEG_FIXED) and ctr[0][1] == 2
    # remove the middle feature, propagate, update the survivors
    bk.clear_feature(1)
    f.clear_feature(1)
    check("after clear_feature")
    for k in range(4, 9):
        t = 0.005 * k
        u = tapes.imu_at(t)
        bk.propagate_state(u, t)
        f.propagate_state(u, t, False)
        for fid in (2, 0):
            zh, _ = f.h_model(abi.FEAT, f.get_state(), fid)
            r = bk.update(abi.FEAT, zh[:2] + 0.3, np.eye(2) * 0.25, id=fid)
            ro = f.update_direct(abi.FEAT, zh[:2] + 0.3, np.eye(2) * 0.25, id=fid)
            assert list(r) == [ro] * B
    check("after updates")
    bk.keep_only_features([0])
    f.keep_only_features([0])
    check("after keep_only_features")


def test_chunked_sharded_and_round1_mappings(pkg, oracle_mod, monkeypatch):
    cfg = _default_cfg()
    NF, B = 8, 11
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=60, seed=23, cam_every=5)
    for i in range(len(ev)):
        if ev[i]["kind"] == abi.EV_PROP:
            ev[i]["sigma"][:6] = [0.05, 0.05, 0.05, 0.002, 0.002, 0.002]
        elif ev[i]["kind"] == abi.EV_UPD and ev[i]["type"] == abi.FEAT:
            ev[i]["sigma"][:2] = 0.3
    bias = [0.02, 0.02, 0.02, 0.002, 0.002, 0.002]
    seed = 0xF4A6
    a = pkg.BatchVIEKF(cfg, B=B, N=NF)
    a.perturb_x0(5)
    ra = a.run_tape(ev, idp, results_first=0, results_count=B, seed=seed, bias_range=bias)
    xa, Pa = a.get_state(), a.get_covariance()
    b = pkg.BatchVIEKF(cfg, B=B, N=NF)
    b.perturb_x0(5)
    d = b.upload_tape(ev, idp, seed=seed, bias_range=bias)
    cuts = [0, 1, 2, 3, 5, 9, 10, len(ev) // 2, len(ev)]  # chunk boundaries off the four-event value-resolution window
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        b.run_dev_tape(d, lo, hi - lo, sync=False)
    b.sync()
    b.free_tape(d)
    assert np.array_equal(b.get_state(), xa) and np.array_equal(b.get_covariance(), Pa)
    parts = []
    for off, cnt in ((0, 3), (3, 8)):
        s = pkg.BatchVIEKF(cfg, B=cnt, N=NF)
        s.set_instance_offset(off)
        s.perturb_x0(5)
        s.run_tape(ev, idp, seed=seed, bias_range=bias)
        parts.append((s.get_state(), s.get_covariance()))
    assert np.array_equal(np.concatenate([p[0] for p in parts]), xa)
    assert np.array_equal(np.concatenate([p[1] for p in parts]), Pa)
    ln, _ = a.get_features()
    for group in (1, 64):  # register-tile executor, two warps per filter of the shared-memory executor
        monkeypatch.setenv("VIEKF_GROUP", str(group))
        c = pkg.BatchVIEKF(cfg, B=B, N=NF)
        c.perturb_x0(5)
        rc = c.run_tape(ev, idp, results_first=0, results_count=B, seed=seed, bias_range=bias)
        assert np.array_equal(ra, rc)
        xc, Pc = c.get_state(), c.get_covariance()
        for i in range(B):
            tapes.live_compare(xa[i], Pa[i], ln[i], xc[i], Pc[i], ln[i], REL_TOL, f"fragment vs VIEKF_GROUP={group} inst {i}")
