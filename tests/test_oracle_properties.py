"""Pins the CPU oracle with the reference's own tests for this path: the property tests of
reference test/jac_test.cpp:245-487 and the known-answer geometry cases of lib/geometry/src/test.cpp.
The reference seeds from the wall clock (jac_test.cpp:490); here the seeds are fixed."""
import numpy as np
import pytest

from tests.common import abi, jac_test_filter

NF = 4
ITERS = 20


def _blocks(NF):
    b = {"dxPOS": (0, 3), "dxVEL": (3, 3), "dxATT": (6, 3), "dxB_A": (9, 3), "dxB_G": (12, 3), "dxMU": (15, 1), "uA": (0, 3), "uG": (3, 3)}
    for i in range(NF):
        b[f"dxZETA_{i}"] = (16 + 3 * i, 2)
        b[f"dxRHO_{i}"] = (16 + 3 * i + 2, 1)
    return b


def _check_block(row, col, a, fd, tol, NF=NF):
    b = _blocks(NF)
    r0, rn = b[row]
    c0, cn = b[col]
    err = np.abs(a[r0:r0 + rn, c0:c0 + cn] - fd[r0:r0 + rn, c0:c0 + cn]).max()
    assert err <= tol, f"Jacobian block ({row},{col}) err {err}"


# ---- geometry known answers: lib/geometry/src/test.cpp:45-73,154-181 -------------------------------
def test_geometry_known_transform(oracle_mod):
    q = oracle_mod.quat_op(17, [0, 0, 1], [np.pi / 4])
    t = np.array([1.0, 1.0, 0.0])
    # inverse: (-q.rotp(t), q^-1) (xform.h:237-240) == known inverse (test.cpp:45-53)
    ti = -oracle_mod.quat_op(1, q, t, 3)
    qi = oracle_mod.quat_op(12, q)
    q_known = oracle_mod.quat_op(17, [0, 0, 1], [-np.pi / 4])
    assert np.allclose(ti, [-np.sqrt(2), 0, 0], atol=1e-8)
    assert np.allclose(qi * np.sign(qi[0]) * np.sign(q_known[0]), q_known, atol=1e-8)
    # passive transform of p1 (test.cpp:55-63): q.rotp(p - t)
    p2 = oracle_mod.quat_op(1, q, np.array([1.0, 0, 0]) - t, 3)
    assert np.allclose(p2, [-np.sqrt(0.5), -np.sqrt(0.5), 0], atol=1e-8)
    # active transform (test.cpp:65-73): q.rota(p) + t
    p3 = oracle_mod.quat_op(0, q, [1.0, 0, 0], 3) + t
    assert np.allclose(p3, [1 + np.sqrt(0.5), 1 + np.sqrt(0.5), 0], atol=1e-8)


def test_geometry_rotation_direction(oracle_mod):
    # lib/geometry/src/test.cpp:154-181
    q = oracle_mod.quat_op(17, [1, 0, 0], [45.0 * np.pi / 180.0])
    v = np.array([0, 0, 1.0])
    R_true = np.array([[1, 0, 0], [0, 0.70710678118654757, 0.70710678118654757], [0, -0.70710678118654757, 0.70710678118654757]])
    R = oracle_mod.quat_op(2, q, n_out=9).reshape(3, 3)
    v_act = np.array([0, -np.sqrt(0.5), np.sqrt(0.5)])
    v_pas = np.array([0, np.sqrt(0.5), np.sqrt(0.5)])
    assert np.allclose(R, R_true, atol=1e-8)
    assert np.allclose(R.T @ v, v_act, atol=1e-8)
    assert np.allclose(oracle_mod.quat_op(0, q, v, 3), v_act, atol=1e-8)
    assert np.allclose(oracle_mod.quat_op(1, q, v, 3), v_pas, atol=1e-8)
    assert np.allclose(R @ v, v_pas, atol=1e-8)


def test_geometry_properties(oracle_mod):
    # exp/log, boxplus/boxminus, from_R, from_two_unit_vectors, euler round trips (test.cpp:183-320)
    rng = np.random.default_rng(3)
    for _ in range(100):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        w = rng.uniform(-1, 1, 3)
        qe = oracle_mod.quat_op(3, q, w)
        assert np.allclose(oracle_mod.quat_op(4, qe, n_out=3), w, atol=1e-8)
        q2 = oracle_mod.quat_op(6, q, w)
        assert np.allclose(oracle_mod.quat_op(7, q2, q, 3), w, atol=1e-8)
        R = oracle_mod.quat_op(2, q, n_out=9)
        qr = oracle_mod.quat_op(11, R)
        assert np.allclose(qr * np.sign(qr[0]) * np.sign(q[0]), q, atol=1e-8)
        v = rng.normal(size=3)
        assert np.allclose(R.reshape(3, 3) @ v, oracle_mod.quat_op(1, q, v, 3), atol=1e-10)
        assert np.allclose(R.reshape(3, 3).T @ v, oracle_mod.quat_op(0, q, v, 3), atol=1e-10)
        u1 = rng.normal(size=3)
        u1 /= np.linalg.norm(u1)
        u2 = rng.normal(size=3)
        u2 /= np.linalg.norm(u2)
        q12 = oracle_mod.quat_op(8, u1, u2)
        assert np.allclose(oracle_mod.quat_op(0, q12, u1, 3), u2, atol=1e-8)  # rota(u) = v (test.cpp:262-277)
        rpy = rng.uniform(-1.4, 1.4, 3)
        assert np.allclose(oracle_mod.quat_op(10, oracle_mod.quat_op(9, rpy), n_out=3), rpy, atol=1e-8)


# ---- VI_EKF.manifold: test/jac_test.cpp:245-280 -----------------------------------------------------
def test_manifold(oracle_mod):
    rng = np.random.default_rng(11)
    for _ in range(ITERS):
        f, x, u = jac_test_filter(oracle_mod, rng, NF)
        _, x2, _ = jac_test_filter(oracle_mod, rng, NF)
        n = f.MAX_DX
        x3 = f.boxplus(x, np.zeros(n))
        assert np.allclose(x3, x, atol=1e-8)
        dx2 = f.boxminus(x2, x)
        x3 = f.boxplus(x, dx2)
        assert np.allclose(x3[:6], x2[:6], atol=1e-8)
        s = np.sign(x3[6]) * np.sign(x2[6])
        assert np.allclose(s * x3[6:10], x2[6:10], atol=1e-8)
        assert np.allclose(x3[10:17], x2[10:17], atol=1e-8)
        for i in range(NF):
            z3 = oracle_mod.quat_op(16, x3[17 + 5 * i: 21 + 5 * i], n_out=3)
            z2 = oracle_mod.quat_op(16, x2[17 + 5 * i: 21 + 5 * i], n_out=3)
            assert np.allclose(z3, z2, atol=1e-8)
            assert abs(x3[21 + 5 * i] - x2[21 + 5 * i]) < 1e-8
        dx = rng.uniform(-1, 1, n)
        x3 = f.boxplus(x, dx)
        assert np.allclose(f.boxminus(x3, x), dx, atol=1e-8)
        dx1, dx2 = rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)
        d = f.boxminus(f.boxplus(x, dx1), f.boxplus(x, dx2))
        assert np.linalg.norm(d) <= np.linalg.norm(d - dx2) + 1e-12  # literal (jac_test.cpp:276)


def _f_tilde(f, x_tilde, x_hat, u, dt):
    # test/jac_test.cpp:283-304
    x = f.boxplus(x_hat, x_tilde)
    dx, _, _ = f.dynamics(x, u)
    dxh, _, _ = f.dynamics(x_hat, u)
    xp, xm = f.boxplus(x, dx * dt), f.boxplus(x, -dx * dt)
    xhp, xhm = f.boxplus(x_hat, dxh * dt), f.boxplus(x_hat, -dxh * dt)
    return (f.boxminus(xp, xhp) - f.boxminus(xm, xhm)) / (2 * dt)


# ---- VI_EKF.dfdx_test / dfdu_test: test/jac_test.cpp:306-415 ------------------------------------------
def test_dfdx_dfdu(oracle_mod):
    rng = np.random.default_rng(5)
    # eps, dt large enough that the perturbations clear the 1e-8 switches of Quat::log (quat.h:146-162) and
    # q_feat_boxminus (math_helper.h:31) and stay out of acos' ill-conditioned range near 1
    eps, dt = 1e-3, 1e-2
    for _ in range(6):
        f, x_hat, u = jac_test_filter(oracle_mod, rng, NF)
        n = f.MAX_DX
        _, A, G = f.dynamics(x_hat, u)
        fdA = np.zeros((n, n))
        for i in range(n):
            e = np.zeros(n)
            e[i] = eps
            fdA[:, i] = (_f_tilde(f, e, x_hat, u, dt) - _f_tilde(f, -e, x_hat, u, dt)) / (2 * eps)
        for r, c in [("dxPOS", "dxVEL"), ("dxPOS", "dxATT"), ("dxVEL", "dxVEL"), ("dxVEL", "dxATT"), ("dxVEL", "dxB_A"),
                     ("dxVEL", "dxB_G"), ("dxVEL", "dxMU"), ("dxATT", "dxATT"), ("dxATT", "dxB_G")]:
            _check_block(r, c, A, fdA, 1e-2)
        for i in range(NF):
            zk, rk = f"dxZETA_{i}", f"dxRHO_{i}"
            for r, c in [(zk, "dxVEL"), (zk, "dxB_G"), (zk, zk), (zk, rk), (rk, "dxVEL"), (rk, "dxB_G"), (rk, zk), (rk, rk)]:
                _check_block(r, c, A, fdA, 5e-1)
        # everything else in A must be structurally zero (SURVEY 7 "structure to exploit")
        mask = np.zeros((n, n), bool)
        b = _blocks(NF)
        def mark(r, c):
            mask[b[r][0]:b[r][0] + b[r][1], b[c][0]:b[c][0] + b[c][1]] = True
        for r, c in [("dxPOS", "dxVEL"), ("dxPOS", "dxATT"), ("dxVEL", "dxVEL"), ("dxVEL", "dxATT"), ("dxVEL", "dxB_A"),
                     ("dxVEL", "dxB_G"), ("dxVEL", "dxMU"), ("dxATT", "dxATT"), ("dxATT", "dxB_G")]:
            mark(r, c)
        for i in range(NF):
            zk, rk = f"dxZETA_{i}", f"dxRHO_{i}"
            for r in (zk, rk):
                for c in ("dxVEL", "dxB_G", zk, rk):
                    mark(r, c)
        assert np.all(A[~mask] == 0.0)
        # dfdu: perturbing the biases == perturbing the inputs (jac_test.cpp:373-415)
        fdG = np.zeros((n, 6))
        for i in range(6):
            e = np.zeros(n)
            e[9 + i] = eps
            fdG[:, i] = (_f_tilde(f, e, x_hat, u, dt) - _f_tilde(f, -e, x_hat, u, dt)) / (2 * eps)
        _check_block("dxVEL", "uA", G, fdG, 1e-2)
        _check_block("dxVEL", "uG", G, fdG, 1e-2)
        _check_block("dxATT", "uG", G, fdG, 1e-2)
        for i in range(NF):
            _check_block(f"dxZETA_{i}", "uG", G, fdG, 5e-1)
            _check_block(f"dxRHO_{i}", "uG", G, fdG, 5e-1)


# ---- VI_EKF.h_test: test/jac_test.cpp:172-226,417-443 -------------------------------------------------
def _htest(oracle_mod, f, mtype, id, dim, tol=1e-3):
    x0 = f.get_state()
    z0, H = f.h_model(mtype, x0, id)
    a = H[:dim, :]
    n = f.MAX_DX
    fd = np.zeros((dim, n))
    eps = 1e-6
    for i in range(n):
        e = np.zeros(n)
        e[i] = eps
        zp, _ = f.h_model(mtype, f.boxplus(x0, e), id)
        if mtype == abi.QZETA:
            fd[:, i] = oracle_mod.quat_op(14, zp, z0, 2) / eps
        elif mtype == abi.ATT:
            fd[:, i] = oracle_mod.quat_op(7, zp, z0, 3) / eps
        else:
            fd[:, i] = (zp[:dim] - z0[:dim]) / eps
    thr = max(tol * np.linalg.norm(a), tol)
    assert np.abs(a - fd).max() <= thr, f"H of type {mtype} id {id}: err {np.abs(a - fd).max()} thr {thr}"


@pytest.mark.parametrize("drag", [True, False])
def test_h_models(oracle_mod, drag):
    rng = np.random.default_rng(17)
    for _ in range(8):
        f, x0, u0 = jac_test_filter(oracle_mod, rng, NF, drag=drag)
        _htest(oracle_mod, f, abi.ACC, 0, 2 if drag else 3)
        _htest(oracle_mod, f, abi.POS, 0, 3)
        _htest(oracle_mod, f, abi.VEL, 0, 3)
        _htest(oracle_mod, f, abi.ALT, 0, 1)
        _htest(oracle_mod, f, abi.ATT, 0, 3)
        for i in range(NF):
            _htest(oracle_mod, f, abi.FEAT, i, 2, 1e-1)
            _htest(oracle_mod, f, abi.QZETA, i, 2)
            _htest(oracle_mod, f, abi.DEPTH, i, 1)
            _htest(oracle_mod, f, abi.INV_DEPTH, i, 1)


# ---- VI_EKF.KF_reset_test: test/jac_test.cpp:446-487 ---------------------------------------------------
def test_keyframe_reset(oracle_mod):
    rng = np.random.default_rng(23)
    for _ in range(ITERS):
        f, xm, u0 = jac_test_filter(oracle_mod, rng, NF)
        xp, N = f.keyframe_reset_jac(xm)
        em, ep = oracle_mod.quat_op(10, xm[6:10], n_out=3), oracle_mod.quat_op(10, xp[6:10], n_out=3)
        assert abs(em[0] - ep[0]) < 1e-8 and abs(em[1] - ep[1]) < 1e-8 and abs(ep[2]) < 1e-8
        n = f.MAX_DX
        fd = np.zeros((n, n))
        eps = 1e-6
        for i in range(n):
            e = np.zeros(n)
            e[i] = eps
            xpp, _ = f.keyframe_reset_jac(f.boxplus(xm, e))
            fd[:, i] = f.boxminus(xpp, xp) / eps
        _check_block("dxPOS", "dxPOS", N, fd, 1e-3)
        _check_block("dxATT", "dxATT", N, fd, 1e-1)
    assert np.abs(N - fd).max() <= 1e-1


# ---- feature bookkeeping semantics (vi_ekf_feat.cpp) ---------------------------
def test_feature_bookkeeping(oracle_mod):
    cfg = abi.default_config()
    cfg.use_keyframe_reset = 0
    f = oracle_mod.OracleVIEKF(cfg, 3)
    f.propagate_state(np.zeros(6), 0.0)
    R = np.eye(2) * 0.25
    assert f.add_measurement(0.01, [300.0, 200.0], abi.FEAT, R, True, 7) == abi.MEAS_NEW_FEATURE
    assert f.tracked_features() == [0]  # id assigned internally, caller id 7 ignored (Q1)
    for k in range(2):
        assert f.add_measurement(0.01, [310.0 + k, 210.0], abi.FEAT, R, True, 100 + k) == abi.MEAS_NEW_FEATURE
    assert f.tracked_features() == [0, 1, 2]
    # full: init refused but add_measurement still reports NEW_FEATURE (Q2)
    assert f.add_measurement(0.01, [100.0, 100.0], abi.FEAT, R, True, 55) == abi.MEAS_NEW_FEATURE
    assert f.get_len_features() == 3
    x = f.get_state()
    P = f.get_covariance()
    f.clear_feature(1)
    assert f.tracked_features() == [0, 2]
    x2, P2 = f.get_state(), f.get_covariance()
    assert np.array_equal(x2[22:27], x[27:32]) and np.all(x2[27:] == 0)
    keep = [i for i in range(25) if not (19 <= i < 22)]
    assert np.array_equal(P2[:22, :22], P[np.ix_(keep, keep)])
    assert np.all(P2[22:, :] == 0) and np.all(P2[:, 22:] == 0)
    f.keep_only_features([2])
    assert f.tracked_features() == [2]
    assert f.init_feature([320.0, 240.0], 0, float("nan"))
    assert f.tracked_features() == [2, 3]
    assert abs(f.get_state()[17 + 5 + 4] - 1.0 / (2 * 1.5)) < 1e-15  # NaN depth -> 1/(2 min_depth) (vi_ekf_feat.cpp:22-26)
