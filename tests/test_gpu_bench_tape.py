"""Parity of the engine with the CPU oracle on the tape bench.py times (sim.bench_workload: waypoint flight, Philox IMU / pixel
noise sigma 1.0 / 0.02 / 0.5 px, per-instance IMU bias, perturbed x0 -- a workload that lives near the Mahalanobis gate),
through the kernel instantiations that carry the reported numbers:

  configs[3]  N = 25  viekf_tile_kernel<3,3>   1000 IMU steps
  configs[4]  N = 35  viekf_tile_kernel<6,2>    200 IMU steps
  configs[2]  N = 8   viekf_frag_kernel        1000 IMU steps   (and viekf_tile_kernel<1,6>, the round-1 mapping)
  configs[1]  N = 0   viekf_octet_kernel       1000 IMU steps

Per-event meas_result_t, feature ids, live counts, flags and counters bit-exact; x / P within 1e-9 relative (north_star).
Instances whose closest gate decision sits within 1e-7 of the threshold are excluded and counted (never re-seeded).
"""
import os

import numpy as np
import pytest

from tests.common import REL_TOL, abi
from tests import tapes
from tests.test_gpu_parity import _stats_close

pytestmark = pytest.mark.gpu

SOAK = int(os.environ.get("VIEKF_SOAK", "1"))  # VIEKF_SOAK=k: k times the IMU steps and k times the oracle sample (evidence runs, profiles/)
CASES = [  # N, B, IMU steps, oracle sample, forced filters per CTA of the tile executor, forced VIEKF_GROUP (None = the engine's choice)
    pytest.param(25, 512, 1000 * SOAK, 64 * SOAK, 3, None, id="configs3_N25_tile33"),
    pytest.param(35, 304, 200 * SOAK, 32 * SOAK, 2, None, id="configs4_N35_tile62"),
    pytest.param(8, 1024, 1000 * SOAK, 64 * SOAK, None, None, id="configs2_N8_frag"),
    pytest.param(8, 1024, 1000 * SOAK, 64 * SOAK, 6, 1, id="N8_tile16"),
    pytest.param(4, 1024, 1000 * SOAK, 64 * SOAK, None, None, id="N4_frag_live_block_guards"),
    pytest.param(0, 1024, 1000 * SOAK, 64 * SOAK, None, None, id="configs1_N0_octet"),
]


def _run_oracle(oracle_mod, sim, cfg, N, ev, idp, nb, count):
    return oracle_mod.run_batch(cfg, N, ev, idp, count=count, nthreads=32, seed=sim.BENCH_SEED, bias_range=sim.BENCH_BIAS,
                                perturb_seed=sim.BENCH_SEED + 1, n_buckets=nb)


@pytest.mark.parametrize("N,B,steps,sample,fpc,group", CASES)
def test_bench_tape_parity(pkg, oracle_mod, monkeypatch, N, B, steps, sample, fpc, group):
    sim = pkg.sim
    if fpc is not None:
        monkeypatch.setenv("VIEKF_TILE_FPC", str(fpc))
    if group is not None:
        monkeypatch.setenv("VIEKF_GROUP", str(group))
    n_slices = steps // sim.BENCH_IMU_PER_STEP
    cfg, ev, idp, info, slices = sim.bench_workload(N, n_slices)
    nb = info["n_buckets"]
    ref = _run_oracle(oracle_mod, sim, cfg, N, ev, idp, nb, sample)
    # one launch over the whole tape: per-event results of the sampled instances
    bk = pkg.BatchVIEKF(cfg, B=B, N=N)
    bk.perturb_x0(sim.BENCH_SEED + 1)
    res = bk.run_tape(ev, idp, results_first=0, results_count=sample, seed=sim.BENCH_SEED, bias_range=sim.BENCH_BIAS)
    ok = ref["margins"][:, 0] > 1e-7
    excluded = int((~ok).sum())
    print(f"N={N}: {excluded} of {sample} instances excluded for a gate margin < 1e-7; gated updates per filter: median "
          f"{np.median(ref['flags'][:, 1]):.0f}, max {ref['flags'][:, 1].max()}")
    assert excluded <= sample // 8
    idx = [b for b in range(sample) if ok[b]]
    assert np.array_equal(res[idx], ref["results"][idx]), "meas_result_t per event differ on the bench tape"
    x, P = bk.get_state(0, sample), bk.get_covariance(0, sample)
    ln, ids = bk.get_features(0, sample)
    fl, ctr = bk.get_flags(0, sample)
    worst = (0.0, 0.0)
    # A filter that has DIVERGED (more than 40 % of its updates gated: it no longer tracks, its correlation matrix is 20-30 x worse
    # conditioned than a healthy one's and amplifies the last-bit differences of the two evaluation orders) keeps bit-exact decisions,
    # ids, flags and counters but is held to 1e-7 on x / P; such filters are counted and printed (none at the default sizes).
    n_upd = int(sum(1 for i in range(len(ev)) if ev[i]["kind"] == abi.EV_UPD))
    diverged = [b for b in idx if ref["flags"][b, 1] > 0.4 * n_upd]
    for b in idx:
        assert ln[b] == ref["len"][b] and list(ids[b][: ln[b]]) == list(ref["ids"][b][: ln[b]]), f"inst {b}: features differ"
        assert fl[b] == ref["flags"][b, 0] and list(ctr[b]) == list(ref["flags"][b, 1:4]), f"inst {b}: flags / counters differ"
        tol = 1e-7 if b in diverged else REL_TOL
        ex, eP = tapes.live_compare(x[b], P[b], ln[b], ref["x"][b], ref["P"][b], ref["len"][b], tol, f"bench tape N={N} inst {b}")
        if b not in diverged:
            worst = (max(worst[0], ex), max(worst[1], eP))
    print(f"N={N}, {steps} IMU steps: worst state rel err {worst[0]:.2e}, covariance rel err {worst[1]:.2e} over {len(idx) - len(diverged)} "
          f"tracking filters; {len(diverged)} diverged filter(s) (> 40 % of {n_upd} updates gated) held to 1e-7")
    assert len(diverged) <= max(1, sample // 8)
    if N > 0:
        heavy = int(np.argmax(ref["flags"][:, 1]))
        assert ref["flags"][heavy, 1] >= 20, "the sample should contain a filter that gates often"
    # the bench's own slicing (one launch per slice of the resident tape) is bit-identical to the single launch
    b2 = pkg.BatchVIEKF(cfg, B=B, N=N)
    b2.perturb_x0(sim.BENCH_SEED + 1)
    d = b2.upload_tape(ev, idp, seed=sim.BENCH_SEED, bias_range=sim.BENCH_BIAS)
    for (s, e) in slices:
        b2.run_dev_tape(d, s, e - s, sync=False)
    b2.sync()
    b2.free_tape(d)
    assert np.array_equal(b2.get_state(0, sample), x) and np.array_equal(b2.get_covariance(0, sample), P)
    fl_all, _ = b2.get_flags()
    assert (fl_all & abi.FLAG_NAN_STATE).sum() == (bk.get_flags()[0] & abi.FLAG_NAN_STATE).sum()


def test_bench_tape_statistics(pkg, oracle_mod):
    """The Monte-Carlo statistics block (the only thing NCCL reduces) over 1,024 instances of the bench tape equals the
    oracle's: counts, error sums, NEES over the filters where it is defined, NaN-flagged filters; NEES16 is of order 16."""
    sim = pkg.sim
    N, B, n_slices = 25, 1024, 5
    cfg, ev, idp, info, slices = sim.bench_workload(N, n_slices)
    nb = info["n_buckets"]
    ref = _run_oracle(oracle_mod, sim, cfg, N, ev, idp, nb, B)
    bk = pkg.BatchVIEKF(cfg, B=B, N=N)
    bk.perturb_x0(sim.BENCH_SEED + 1)
    bk.stats_reset(nb)
    bk.run_tape(ev, idp, seed=sim.BENCH_SEED, bias_range=sim.BENCH_BIAS)
    st = bk.stats(nb)
    _stats_close(st, ref["stats"], rtol=1e-6)
    assert np.array_equal(st[:, [0, 37, 38, 39]], ref["stats"][:, [0, 37, 38, 39]])
    assert np.all(st[:, 0] + st[:, 38] == B)
    nees16 = st[:, 33] / np.maximum(st[:, 37], 1)
    assert np.all(np.isfinite(st)) and 8.0 < nees16[0] < 40.0, f"NEES16 {nees16}"
    fl, ctr = bk.get_flags()
    assert np.array_equal(fl, ref["flags"][:, 0]) and np.array_equal(ctr, ref["flags"][:, 1:4])


def test_statistics_across_keyframe_reset(pkg, oracle_mod):
    """STAT compares like with like after keyframe_reset (vi_ekf_kfr.cpp:56-157): on a noise-free tape with resets the error
    against the global truth stays ~0 (it was the node offset before), and engine == oracle."""
    sim = pkg.sim
    N, B = 12, 4
    cfg = abi.default_config()
    cfg.keyframe_overlap_threshold = 0.95  # resets on nearly every frame with a dropped feature
    truth = sim.simulate(cfg, N, n_imu=400)
    ev, idp, info = sim.build_tape(truth, cfg, N, use_alt=False, stat_every=20, depth_known=True, imu_sigma=(0,) * 6, pix_sigma=0.0)
    nb = info["n_buckets"]
    ref = oracle_mod.run_batch(cfg, N, ev, idp, count=B, nthreads=4, n_buckets=nb)
    bk = pkg.BatchVIEKF(cfg, B=B, N=N)
    bk.stats_reset(nb)
    bk.run_tape(ev, idp)
    st = bk.stats(nb)
    # counts and error sums; the NEES columns are not compared here: right after a reset the pose block of P is (nearly) singular
    assert np.array_equal(st[:, [0, 37, 38, 39]], ref["stats"][:, [0, 37, 38, 39]])
    assert np.allclose(st[:, 1:33], ref["stats"][:, 1:33], rtol=1e-6, atol=1e-10)
    pose, _ = bk.get_global_pose()
    assert np.abs(pose[0, :3]).max() > 0.05, "the tape should have moved the keyframe node"
    rmse_pos = np.sqrt(st[:, 17:20] / np.maximum(st[:, 0:1], 1))
    rmse_yaw = np.sqrt(st[:, 25] / np.maximum(st[:, 0], 1))
    assert rmse_pos.max() < 0.05 and rmse_yaw.max() < 0.02, f"rmse_pos {rmse_pos.max()}, rmse_yaw {rmse_yaw.max()}"


def test_reference_controller_stream_parity(pkg, oracle_mod):
    """The same comparison on streams flown by the reference's own closed loop (refsim.py: NLC + PID + waypoint manager,
    lib/multirotor_sim/src/controller.cpp:18-62,229-279), with the bench's noise model: engine == oracle."""
    sim = pkg.sim
    N, B, sample = 12, 48, 32
    cfg = abi.default_config()
    cfg.use_keyframe_reset = 0
    truth = sim.simulate(cfg, N, n_imu=300, controller="reference")
    ev, idp, info = sim.build_tape(truth, cfg, N, stat_every=20, depth_known=True)
    nb = info["n_buckets"]
    ref = _run_oracle(oracle_mod, sim, cfg, N, ev, idp, nb, sample)
    bk = pkg.BatchVIEKF(cfg, B=B, N=N)
    bk.stats_reset(nb)
    bk.perturb_x0(sim.BENCH_SEED + 1)
    res = bk.run_tape(ev, idp, results_first=0, results_count=sample, seed=sim.BENCH_SEED, bias_range=sim.BENCH_BIAS)
    ok = ref["margins"][:, 0] > 1e-7
    assert (~ok).sum() <= sample // 8
    idx = [b for b in range(sample) if ok[b]]
    assert np.array_equal(res[idx], ref["results"][idx])
    x, P = bk.get_state(0, sample), bk.get_covariance(0, sample)
    ln, ids = bk.get_features(0, sample)
    for b in idx:
        assert ln[b] == ref["len"][b] and list(ids[b][: ln[b]]) == list(ref["ids"][b][: ln[b]])
        tapes.live_compare(x[b], P[b], ln[b], ref["x"][b], ref["P"][b], ref["len"][b], REL_TOL, f"reference-controller stream inst {b}")
