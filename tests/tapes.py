"""Seeded event tapes for the parity tests (TEST INFRASTRUCTURE).

`consistent_tape` drives one CPU-oracle filter while it writes the tape, so every measurement is the oracle's own
prediction plus noise (most pass the Mahalanobis gate, a few deliberate outliers do not) and the tape exercises
propagate, every measurement model, feature init / keep / clear churn, keyframe resets and statistics events --
the call pattern of the reference's ROS image / IMU callbacks (src/vi_ekf_ros.cpp:147-232,254-329).
"""
import numpy as np

import __graft_entry__ as g

pkg = g.load_package()
abi = pkg.abi


def imu_at(t, rng=None, noise=0.0):
    """Smooth synthetic IMU sample in the IMU frame (hover-ish: -g on z plus slow oscillations)."""
    acc = np.array([0.4 * np.sin(1.3 * t), 0.3 * np.cos(0.9 * t), -9.80665 + 0.5 * np.sin(0.7 * t)])
    gyr = np.array([0.15 * np.sin(1.1 * t), 0.12 * np.cos(0.8 * t), 0.1 * np.sin(0.5 * t)])
    u = np.concatenate([acc, gyr])
    if rng is not None and noise > 0:
        u = u + rng.normal(size=6) * noise * np.array([1, 1, 1, 0.02, 0.02, 0.02])
    return u


def consistent_tape(oracle_mod, cfg, NF, steps=60, seed=0, dt=0.005, cam_every=6, churn_every=2, extra_models=True,
                    outlier_every=17, stat_every=10, t0=0.0, depth_known=True):
    """Returns (events, id_pool, oracle_filter_after_tape, n_stat_buckets)."""
    rng = np.random.default_rng(seed)
    f = oracle_mod.OracleVIEKF(cfg, NF, hist=2)
    tb = pkg.TapeBuilder()
    next_id = 0  # caller-side id counter; equals the filter's next_feature_id_ while no init is refused
    t = t0
    bucket = 0
    frame = 0

    def upd(mtype, z, R, id=-1, depth=float("nan"), active=True):
        tb.update(mtype, z, R, active=active, id=id, depth=depth)
        return f.update_direct(mtype, z, R, active=active, id=id, depth=depth)

    u = imu_at(t)
    tb.propagate(u, t)
    f.propagate_state(u, t, False)  # latches the start time
    for k in range(1, steps + 1):
        t = t0 + k * dt
        u = imu_at(t, rng, 0.05)
        tb.propagate(u, t)
        f.propagate_state(u, t, False)
        x = f.get_state()
        # accelerometer every IMU sample (src/vi_ekf_ros.cpp:180-191)
        if cfg.use_drag_term:
            zh, _ = f.h_model(abi.ACC, x)
            z = zh[:2] + rng.normal(size=2) * 0.3
            if outlier_every and k % outlier_every == 0:
                z = z + 40.0
            upd(abi.ACC, z, np.eye(2))
        else:
            zh, _ = f.h_model(abi.ACC, x)
            upd(abi.ACC, zh[:3] + rng.normal(size=3) * 0.3, np.eye(3))
        if k % 8 == 0:
            zh, _ = f.h_model(abi.ALT, f.get_state())
            upd(abi.ALT, zh[:1] + rng.normal(size=1) * 0.05, np.array([[0.1]]))
        if extra_models and k % 11 == 0:
            x = f.get_state()
            upd(abi.POS, x[0:3] + rng.normal(size=3) * 0.05, np.eye(3) * 0.1)
            x = f.get_state()
            upd(abi.VEL, x[3:6] + rng.normal(size=3) * 0.05, np.eye(3) * 0.1)
            x = f.get_state()
            qz = oracle_mod.quat_op(6, x[6:10], rng.normal(size=3) * 0.02)
            upd(abi.ATT, qz, np.eye(3) * 0.01)
            upd(abi.POS, x[0:3], np.eye(3) * 0.1, active=False)  # inactive: logged only (vi_ekf_meas.cpp:230)
        if NF > 0 and k % cam_every == 0:
            frame += 1
            tracked = f.tracked_features()
            keep = list(tracked)
            if churn_every and frame % churn_every == 0 and len(keep) > 0:
                keep.pop(int(rng.integers(0, len(keep))))
            tb.keep_only_features(keep + [10_000 + frame])  # an id the filter never tracked is simply ignored
            f.keep_only_features(keep + [10_000 + frame])
            # reverse order, as the LIFO tie rule of the measurement queue applies them
            for fid in reversed(keep):
                zh, _ = f.h_model(abi.FEAT, f.get_state(), fid)
                z = zh[:2] + rng.normal(size=2) * 0.5
                if outlier_every and (frame + fid) % outlier_every == 0:
                    z = z + 60.0
                upd(abi.FEAT, z, np.eye(2) * 0.25, id=fid)
                if extra_models and (frame + fid) % 5 == 0:
                    zh, _ = f.h_model(abi.DEPTH, f.get_state(), fid)
                    upd(abi.DEPTH, zh[:1] * (1 + rng.normal() * 0.02), np.array([[0.5]]), id=fid)
                if extra_models and (frame + fid) % 7 == 0:
                    zh, _ = f.h_model(abi.INV_DEPTH, f.get_state(), fid)
                    upd(abi.INV_DEPTH, zh[:1] + rng.normal() * 0.01, np.array([[0.05]]), id=fid)
                if extra_models and (frame + fid) % 9 == 0:
                    x = f.get_state()
                    loc = f.tracked_features().index(fid)
                    qz = oracle_mod.quat_op(13, x[17 + 5 * loc:17 + 5 * loc + 4], rng.normal(size=2) * 0.005)
                    upd(abi.QZETA, qz, np.eye(2) * 0.01, id=fid)
            # new features up to the capacity, plus one more that the full filter must refuse
            n_new = NF - f.get_len_features()
            for _ in range(n_new):
                pix = np.array([rng.uniform(40, 600), rng.uniform(40, 440)])
                depth = float(rng.uniform(2.0, 12.0)) if (depth_known and rng.uniform() < 0.7) else float("nan")
                r = upd(abi.FEAT, pix, np.eye(2) * 0.25, id=next_id, depth=depth)
                assert r == abi.MEAS_NEW_FEATURE
                next_id += 1
            if frame % 3 == 0:
                r = upd(abi.FEAT, np.array([320.0, 240.0]), np.eye(2) * 0.25, id=next_id + 500)
                assert r == abi.MEAS_NEW_FEATURE and f.get_len_features() == NF
            if extra_models and frame % 4 == 0:
                upd(abi.DEPTH, np.array([3.0]), np.array([[0.5]]), id=777_777)  # untracked id -> MEAS_INVALID
        if stat_every and k % stat_every == 0:
            x = f.get_state()
            xt = x[:17].copy()
            xt[0:6] += rng.normal(size=6) * 0.02
            xt[6:10] = oracle_mod.quat_op(6, xt[6:10], rng.normal(size=3) * 0.01)
            xt[10:17] += rng.normal(size=7) * 0.01
            tb.stat(bucket, xt)
            bucket += 1
    return tb.events(), tb.id_pool(), f, bucket


def perturbed_values(events, B, seed, scale=1.0):
    """Per-instance value streams: instance 0 replays the tape's own z, the others add small seeded perturbations to
    every PROP / UPD value so that their trajectories (and a few gate decisions) differ."""
    ev = events.copy()
    rng = np.random.default_rng(seed)
    blocks = 0
    for i in range(len(ev)):
        if ev[i]["kind"] in (abi.EV_PROP, abi.EV_UPD) and not (ev[i]["kind"] == abi.EV_UPD and ev[i]["type"] in (abi.ATT, abi.QZETA)):
            ev[i]["noise"] = blocks
            blocks += 1
    vals = np.zeros((B, 8 * max(blocks, 1)))
    for i in range(len(ev)):
        nb = int(ev[i]["noise"])
        if nb < 0:
            continue
        z = ev[i]["z"].copy()
        if ev[i]["kind"] == abi.EV_PROP:
            sc = np.array([0.05, 0.05, 0.05, 0.002, 0.002, 0.002, 0, 0]) * scale
        else:
            sc = np.zeros(8)
            zd = int(ev[i]["zdim"])
            base = {abi.FEAT: 0.3, abi.ACC: 0.1, abi.ALT: 0.02, abi.POS: 0.02, abi.VEL: 0.02, abi.DEPTH: 0.02, abi.INV_DEPTH: 0.002}
            sc[:zd] = base.get(int(ev[i]["type"]), 0.0) * scale
        pert = rng.normal(size=(B, 8)) * sc
        pert[0] = 0.0
        vals[:, 8 * nb:8 * nb + 8] = z[None, :] + pert
    return ev, vals


def live_compare(x_g, P_g, len_g, x_o, P_o, len_o, tol, tag=""):
    """Compare the live region of (x, P) of the engine against the oracle."""
    from tests.common import cov_rel_err, state_rel_err
    assert int(len_g) == int(len_o), f"{tag}: len {len_g} != {len_o}"
    nx, n = 17 + 5 * int(len_o), 16 + 3 * int(len_o)
    ex = state_rel_err(x_g[:nx], x_o[:nx])
    eP = cov_rel_err(P_g[:n, :n], P_o[:n, :n])
    assert ex <= tol, f"{tag}: state rel err {ex:.3e} > {tol}"
    assert eP <= tol, f"{tag}: covariance rel err {eP:.3e} > {tol}"
    return ex, eP
