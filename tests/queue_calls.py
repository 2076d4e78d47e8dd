"""Recorded call logs of the streaming call pattern (propagate_state / add_measurement / handle_measurements) and their replay
on either filter: the oracle (OracleVIEKF) or the engine in queue mode (BatchVIEKF after enable_queue).  TEST INFRASTRUCTURE."""
import numpy as np

CALL_DTYPE = np.dtype([("kind", "<i4"), ("id", "<i4"), ("t", "<f8"), ("v", "<f8", (6,))])  # kind: 0 propagate, 1 add FEAT, 2 handle
FEAT_R = 0.25 * np.eye(2)  # simulator.cpp:201 feat_R_, passed through by test/vi_ekf_test.cpp:31


def config0_calls(pkg, cfg, NF=12, n_imu=300, seed=11):
    """BASELINE configs[0]: the call sequence of the reference's test/vi_ekf_test.cpp:24-33,58 -- imu_cb -> propagate_state(z, t,
    true); feat_cb -> add_measurement(t, pix, FEAT, R_pix, true, id, NAN) per feature, then handle_measurements() -- over the
    restated multirotor_sim stream (250 Hz IMU, 30 Hz camera)."""
    truth = pkg.sim.simulate(cfg, NF, n_imu=n_imu, imu_rate=250.0, cam_rate=30.0, alt_rate=0)
    rng = np.random.default_rng(seed)
    calls = []
    frames = list(truth.frames)
    fi = 0
    for k in range(len(truth.imu_t)):
        t = float(truth.imu_t[k])
        u = truth.imu_u[k] + np.array([0.05] * 3 + [0.002] * 3) * rng.standard_normal(6)
        calls.append((0, -1, t, u))
        while fi < len(frames) and frames[fi][0] <= t + 1e-9:
            tf, vis, _ = frames[fi]
            fi += 1
            for (fid, pix, depth) in vis:
                z = np.zeros(6)
                z[:2] = np.asarray(pix, dtype=np.float64) + 0.5 * rng.standard_normal(2)
                calls.append((1, int(fid), float(tf), z))
            calls.append((2, -1, float(tf), np.zeros(6)))
    return np.array(calls, dtype=CALL_DTYPE)


def replay(calls, filt, abi, engine):
    """Returns (add_measurement results, gated FEAT ids of instance 0 per handle_measurements call, number of updates applied)."""
    adds, gated, n_upd = [], [], 0
    for c in calls:
        if c["kind"] == 0:
            (filt.queue_propagate_state if engine else filt.propagate_state)(c["v"], float(c["t"]), True)
        elif c["kind"] == 1:
            adds.append(int(filt.add_measurement(float(c["t"]), c["v"][:2], abi.FEAT, FEAT_R, True, int(c["id"]), float("nan"))))
        else:
            if engine:
                ty, ids, res = filt.handle_measurements()
                gated.append([int(ids[j]) for j in range(len(ty)) if ty[j] == abi.FEAT and res[j, 0] == abi.MEAS_GATED])
                n_upd += len(ty)
            else:
                gated.append([int(i) for i in filt.handle_measurements()])
    return adds, gated, n_upd
