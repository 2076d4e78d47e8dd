"""Shared helpers for the parity tests: the reference's init_jacobians_test recipe (test/jac_test.cpp:118-170) with a
FIXED seed (the reference seeds from the wall clock, :490), tolerances, comparison utilities."""
import numpy as np

import __graft_entry__ as g

abi = g.load_package().abi

# north_star: "state and covariance must agree to a relative tolerance of 1e-9 over 1000 steps"
REL_TOL = 1e-9


def jac_test_config(rng, abi=abi, drag=True, partial=True, kf_reset=True, overlap=0.0):
    """Random filter configuration of init_jacobians_test (reference test/jac_test.cpp:118-149)."""
    cfg = abi.default_config()
    x0 = np.zeros(17)
    x0[abi.xATT] = 1.0
    x0[abi.xMU] = 0.2
    x0[0:3] += rng.uniform(-1, 1, 3) * 100.0
    x0[3:6] += rng.uniform(-1, 1, 3) * 10.0
    x0[10:13] += rng.uniform(-1, 1, 3) * 1.0
    x0[13:16] += rng.uniform(-1, 1, 3) * 0.5
    x0[16] += rng.uniform(0, 1) * 0.05
    att_delta = rng.uniform(-1, 1, 3)
    cfg.P0[:] = [1.0] * 16
    cfg.Qx[:] = [1.0] * 16
    cfg.lambda_[:] = [1.0] * 16
    cfg.Qu[:] = [1.0] * 6
    cfg.P0_feat[:] = [1.0] * 3
    cfg.Qx_feat[:] = [1.0] * 3
    cfg.lambda_feat[:] = [1.0] * 3
    cfg.cam_center[:] = [320 - 25 + int(rng.integers(0, 50)), 240 - 25 + int(rng.integers(0, 50))]
    cfg.focal_len[:] = [250 + rng.uniform(0, 50), 250 + rng.uniform(0, 50)]
    q = rng.uniform(-1, 1, 4)
    q /= np.linalg.norm(q)
    cfg.q_b_c[:] = list(np.array([0.5, 0.5, 0.5, 0.5]) + q)  # deliberately non-unit
    cfg.p_b_c[:] = list(rng.uniform(-1, 1, 3) * 0.5)
    cfg.q_b_u[:] = [1.0, 0.0, 0.0, 0.0]  # 17-arg init leaves q_b_u identity
    cfg.min_depth = 2.0
    cfg.use_drag_term = int(drag)
    cfg.use_partial_update = int(partial)
    cfg.use_keyframe_reset = int(kf_reset)
    cfg.keyframe_overlap_threshold = overlap
    return cfg, x0, att_delta


def jac_test_filter(oracle_mod, rng, NF, **kw):
    """Returns (oracle filter with NF random features, x0 (MAX_X), u0) as init_jacobians_test does."""
    cfg, x0, att_delta = jac_test_config(rng, **kw)
    x0[6:10] = oracle_mod.quat_op(6, x0[6:10], att_delta)  # q [+] Random()
    cfg.x0[:] = list(x0)
    f = oracle_mod.OracleVIEKF(cfg, NF)
    for i in range(NF):
        l = np.array([float(rng.integers(0, 640)), float(rng.integers(0, 480))])
        depth = 1.0 + rng.uniform(0, 1) * 20.0
        assert f.init_feature(l, i, depth)
    u0 = np.concatenate([rng.uniform(-1, 1, 3), rng.uniform(-1, 1, 3)])
    return f, f.get_state(), u0


def cov_rel_err(P, Pref):
    """max_ij |P - Pref|_ij / sqrt(Pref_ii Pref_jj): covariance error relative to the natural scale of each entry."""
    d = np.sqrt(np.abs(np.diag(Pref)))
    scale = np.outer(d, d)
    scale = np.where(scale > 0, scale, 1.0)
    return float(np.max(np.abs(P - Pref) / scale))


def state_rel_err(x, xref):
    return float(np.max(np.abs(x - xref) / np.maximum(1.0, np.abs(xref))))
