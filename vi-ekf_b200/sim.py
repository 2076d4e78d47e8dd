"""Synthetic multirotor truth + sensor streams and the in-order event tape built from them.

A reduced restatement of the parts of lib/multirotor_sim the reference's sim-driven test uses
(reference test/vi_ekf_test.cpp:12-87): rigid-body dynamics with linear drag integrated by RK4
(lib/multirotor_sim/src/dynamics.cpp:50-106), IMU synthesis (:59-64, src/simulator.cpp:384-403), ground-plane
landmarks seen by a pinhole camera with features replaced as they leave the image (src/environment.cpp:46-78,
src/simulator.cpp:406-488,680-701) and an altimeter (:492-502), flown over the waypoints of params/sim_params.yaml:19-28.
Two vehicles can fly the waypoints: `controller="reduced"` (default; a plain cascaded P controller on a first-order
rate loop: a smooth, dynamically consistent trajectory, the one every committed fixture and bench number was produced with) and
`controller="reference"`: the reference's own closed loop (nonlinear velocity controller, roll / pitch / yaw-rate PID, hover-throttle
observer, waypoint manager, force / torque dynamics) as restated in refsim.py and pinned against the waypoint-arrival times the
reference's test holds.  Everything here is deterministic; the
per-instance sensor noise of a Monte-Carlo batch is drawn on the device (include/viekf_noise.h).

The tape follows the order in which the reference's measurement queue ends up applying things:
an IMU interval that contains a measurement is integrated in two sub-steps with the same IMU sample, new features
are initialised in arrival order, tracked features are updated in reverse arrival order, and keep_only_features runs
before the frame's measurements as in src/vi_ekf_ros.cpp:282-309.
"""
import numpy as np

from . import abi
from .tape import TapeBuilder

G = 9.80665
WAYPOINTS = np.array([[1, 1, -1, 0.0], [-1, -1, -1, -0.7], [-1, 1, -1, 0.0], [1, -1, -1, 0.7],
                      [1, 1, -2, 0.0], [-1, -1, -2, -0.7], [-1, 1, -2, 0.0], [1, -1, -2, 0.7]], dtype=float)  # sim_params.yaml:19-28


# ---- quaternion helpers, conventions of lib/geometry quat.h ([w,x,y,z], rotp = R v passive, rota = R^T v) ----------
def q_rota(q, v):
    b = q[1:]
    t = 2.0 * np.cross(v, b)
    return v - q[0] * t + np.cross(t, b)


def q_rotp(q, v):
    b = q[1:]
    t = 2.0 * np.cross(v, b)
    return v + q[0] * t + np.cross(t, b)


def q_mul(a, b):
    return np.array([a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                     a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


def q_exp(v):
    n = np.linalg.norm(v)
    if n > 1e-4:
        return np.concatenate([[np.cos(n / 2)], np.sin(n / 2) * v / n])
    q = np.concatenate([[1.0], v / 2])
    return q / np.linalg.norm(q)


def q_log(q):
    n = np.linalg.norm(q[1:])
    if n < 1e-8:
        return np.zeros(3)
    return 2.0 * np.arctan2(n, q[0]) * q[1:] / n


def q_from_euler(r, p, y):
    cp, ct, cs = np.cos(r / 2), np.cos(p / 2), np.cos(y / 2)
    sp, st, ss = np.sin(r / 2), np.sin(p / 2), np.sin(y / 2)
    return np.array([cp * ct * cs + sp * st * ss, sp * ct * cs - cp * st * ss, cp * st * cs + sp * ct * ss, cp * ct * ss - sp * st * cs])


def q_yaw(q):
    return np.arctan2(2.0 * (q[0] * q[3] + q[1] * q[2]), 1.0 - 2.0 * (q[2] * q[2] + q[3] * q[3]))


class Truth:
    """Output of simulate(): everything the tape builder needs, all deterministic."""

    def __init__(self):
        self.imu_t, self.imu_u, self.x_true = [], [], []   # x_true: 17-vector in the filter's state layout at each IMU time
        self.frames = []                                    # (t, [(id, pixel(2), depth)], x_true17)
        self.alt = []                                       # (t, z)


def simulate(cfg, N, n_imu=1000, imu_rate=200.0, cam_rate=30.0, alt_rate=25.0, dt=0.0005, drag=0.1, seed=1, speed=1.0,
             img_size=(640.0, 480.0), controller="reduced"):
    """Flies the waypoints and records IMU samples (IMU frame, noise-free), camera frames with up to N tracked
    landmarks, and altimeter samples.  State: p (inertial NED), v (body), q (inertial->body), w (body)."""
    rng = np.random.default_rng(seed)
    q_bu = np.array(cfg.q_b_u[:])
    q_bc = np.array(cfg.q_b_c[:])
    p_bc = np.array(cfg.p_b_c[:])
    fl = np.array(cfg.focal_len[:])
    cc = np.array(cfg.cam_center[:])
    img = np.array(img_size)
    p = np.array([0.0, 0.0, -2.0])
    v = np.zeros(3)
    q = np.array([1.0, 0, 0, 0])
    w = np.zeros(3)
    wp = 0
    thrust = G

    def control(p, v, q, w):
        nonlocal wp
        if np.linalg.norm(WAYPOINTS[wp, :3] - p) < 0.15:
            wp = (wp + 1) % len(WAYPOINTS)
        tgt = WAYPOINTS[wp]
        v_des = 1.5 * (tgt[:3] - p)
        nv = np.linalg.norm(v_des)
        if nv > speed:
            v_des *= speed / nv
        a_des = 2.0 * (v_des - q_rota(q, v))
        a_des = np.clip(a_des, -3.0, 3.0)
        yaw = q_yaw(q)
        cy, sy = np.cos(yaw), np.sin(yaw)
        a_fwd, a_right = cy * a_des[0] + sy * a_des[1], -sy * a_des[0] + cy * a_des[1]
        th_des = np.clip(-a_fwd / G, -0.35, 0.35)
        ph_des = np.clip(a_right / G, -0.35, 0.35)
        dyaw = (tgt[3] - yaw + np.pi) % (2 * np.pi) - np.pi
        q_des = q_from_euler(ph_des, th_des, yaw + np.clip(dyaw, -0.3, 0.3))
        e = q_mul(np.array([q[0], -q[1], -q[2], -q[3]]), q_des)
        if e[0] < 0:
            e = -e
        w_cmd = 6.0 * q_log(e)
        T = (G - a_des[2]) / max(np.cos(ph_des) * np.cos(th_des), 0.5)
        return T, w_cmd

    def f(p, v, q, w, T, w_cmd):
        dp = q_rota(q, v)
        dv = np.array([0, 0, -T]) - drag * v + q_rotp(q, np.array([0, 0, G])) - np.cross(w, v)
        dw = (w_cmd - w) / 0.05
        return dp, dv, w, dw

    def step(p, v, q, w, T, w_cmd, h):
        k1 = f(p, v, q, w, T, w_cmd)
        s2 = (p + k1[0] * h / 2, v + k1[1] * h / 2, q_mul(q, q_exp(k1[2] * h / 2)), w + k1[3] * h / 2)
        k2 = f(*s2, T, w_cmd)
        s3 = (p + k2[0] * h / 2, v + k2[1] * h / 2, q_mul(q, q_exp(k2[2] * h / 2)), w + k2[3] * h / 2)
        k3 = f(*s3, T, w_cmd)
        s4 = (p + k3[0] * h, v + k3[1] * h, q_mul(q, q_exp(k3[2] * h)), w + k3[3] * h)
        k4 = f(*s4, T, w_cmd)
        dpp = (k1[0] + 2 * k2[0] + 2 * k3[0] + k4[0]) * h / 6
        dvv = (k1[1] + 2 * k2[1] + 2 * k3[1] + k4[1]) * h / 6
        dqq = (k1[2] + 2 * k2[2] + 2 * k3[2] + k4[2]) * h / 6
        dww = (k1[3] + 2 * k2[3] + 2 * k3[3] + k4[3]) * h / 6
        qn = q_mul(q, q_exp(dqq))
        return p + dpp, v + dvv, qn / np.linalg.norm(qn), w + dww

    def x17(p, v, q):
        x = np.zeros(17)
        x[0:3], x[3:6], x[6:10] = p, v, q
        x[16] = drag
        return x

    def cam_pose(p, q):
        return p + q_rota(q, p_bc), q_mul(q, q_bc)

    def project(pt, t_c, q_c):
        pc = q_rotp(q_c, pt - t_c)
        if pc[2] <= 1e-3:
            return None
        pix = fl * pc[:2] / pc[2] + cc
        if (pix < 0).any() or (pix > img).any():
            return None
        return pix, float(np.linalg.norm(pc))

    out = Truth()
    imu_every = int(round(1.0 / (imu_rate * dt)))
    veh = con = None
    ft = np.zeros(4)
    if controller == "reference":
        # the reference's vehicle and controller (refsim.py); the controller runs at the IMU rate as in Simulator::run
        # (simulator.cpp:100-121: the time step is the IMU period), the dynamics on the finer grid with the forces held
        from . import refsim
        prm = dict(enable_dynamics_noise=False, drag_constant=drag)
        veh = refsim.Vehicle(refsim.State(p, q, v, w), prm, seed)
        con = refsim.ReferenceController(WAYPOINTS, prm)
    elif controller != "reduced":
        raise ValueError("controller must be 'reduced' or 'reference'")
    cam_every = int(np.ceil(1.0 / (cam_rate * dt) - 1e-9))   # ">= 1/rate" on the 0.5 ms grid (simulator.cpp:409): 67 steps at 30 Hz
    alt_every = int(round(1.0 / (alt_rate * dt))) if alt_rate > 0 else 0
    tracked = []  # [id, point]
    next_id = 0
    k = 0
    n_steps = n_imu * imu_every
    while len(out.imu_t) < n_imu + 1:
        t = k * dt
        if veh is not None:
            if k % imu_every == 0:
                tc = t + imu_every * dt  # the reference's clock is advanced before the controller runs
                con.get_commanded_state(tc)
                ft = veh.low_level(con.compute_control(tc, veh.x), imu_every * dt)
        else:
            T, w_cmd = control(p, v, q, w)
        if k % imu_every == 0:
            if veh is not None:
                acc_b = veh.imu(veh.x, ft)[0]
            else:
                acc_b = np.array([0, 0, -T]) - drag * v  # specific force in the body frame (dynamics.cpp:62, p_b_u = 0)
            out.imu_t.append(t)
            out.imu_u.append(np.concatenate([q_rotp(q_bu, acc_b), q_rotp(q_bu, w)]))
            out.x_true.append(x17(p, v, q))
        if N > 0 and k > 0 and k % cam_every == 0:
            t_c, q_c = cam_pose(p, q)
            vis = []
            keep = []
            for fid, pt in tracked:
                pr = project(pt, t_c, q_c)
                if pr is not None:
                    vis.append((fid, pr[0], pr[1]))
                    keep.append([fid, pt])
            tracked = keep
            tries = 0
            while len(tracked) < N and tries < 20 * N:
                tries += 1
                pix = 0.45 * rng.uniform(-1, 1, 2) * img
                zeta = np.array([pix[0] / fl[0], pix[1] / fl[1], 1.0])
                zeta /= np.linalg.norm(zeta)
                zI = q_rota(q_c, zeta)
                if zI[2] <= 1e-3:
                    continue
                depth = -t_c[2] / zI[2]
                if depth < 0.5 or depth > 100.0:
                    continue
                pt = t_c + depth * zI
                tracked.append([next_id, pt])
                vis.append((next_id, pix + cc, float(depth)))
                next_id += 1
            out.frames.append((t, vis, x17(p, v, q)))
        if alt_every and k > 0 and k % alt_every == 0:
            out.alt.append((t, -p[2]))
        if veh is not None:
            veh.run(dt, ft)
            p, v, q, w = veh.x.p, veh.x.v, veh.x.q, veh.x.w
        else:
            p, v, q, w = step(p, v, q, w, T, w_cmd, dt)
        k += 1
    out.imu_t = np.array(out.imu_t)
    out.imu_u = np.array(out.imu_u)
    out.x_true = np.array(out.x_true)
    return out


def build_tape(truth, cfg, N, use_acc=True, use_alt=True, use_feat=True, feat_R=0.25, acc_R=1.0, alt_R=0.1, stat_every=0,
               imu_sigma=(1.0, 1.0, 1.0, 0.02, 0.02, 0.02), pix_sigma=0.5, alt_sigma=0.1, depth_known=False):
    """In-order tape of the truth run.  Returns (events, id_pool, info): info counts the IMU steps (= filter-steps), the
    events of each kind, the live feature count at every event (for the flop model) and the feature churn."""
    tb = TapeBuilder()
    live = []           # live feature count seen by each event
    tracked = []
    n_init = n_drop = 0
    info = dict(n_imu=0, n_prop=0, n_feat=0, n_acc=0, n_alt=0, n_init=0, n_drop=0, n_keep=0, n_stat=0)
    meas = []  # (t, kind, payload), time-sorted
    if use_feat:
        for (t, vis, xt) in truth.frames:
            meas.append((t, 0, (vis, xt)))
    if use_alt:
        for (t, z) in truth.alt:
            meas.append((t, 1, z))
    meas.sort(key=lambda m: (m[0], m[1]))
    mi = 0
    bucket = 0
    Rf = np.eye(2) * feat_R

    def ev_done():
        live.append(len(tracked))

    if use_alt and cfg.use_keyframe_reset:
        # keyframe_reset re-zeroes the filter's position, z included (vi_ekf_kfr.cpp:63,135), while the altimeter measures the
        # global altitude (src/vi_ekf_ros.cpp:440,476): the two are inconsistent in the reference itself ("No altimeter for
        # now", vi_ekf_kfr.cpp:113) -- every ALT update after the first reset would be gated
        raise ValueError("build_tape: use_alt needs cfg.use_keyframe_reset = 0 (the altimeter measures the global altitude)")
    tb.propagate(truth.imu_u[0], truth.imu_t[0], sigma=imu_sigma, sample=0)  # first call only latches the start time
    ev_done()
    for k in range(1, len(truth.imu_t)):
        tk, uk = truth.imu_t[k], truth.imu_u[k]
        at_tk = False  # a measurement stamped exactly tk has already propagated the filter there
        while mi < len(meas) and meas[mi][0] <= tk + 1e-9:
            tz, kind, pl = meas[mi]
            mi += 1
            if tz < tk - 1e-9:
                # a split sub-step and the closing sub-step of the interval are the same physical IMU sample:
                # `sample=k` makes every instance draw the same generated noise for both
                tb.propagate(uk, tz, sigma=imu_sigma, sample=k)
                info["n_prop"] += 1
                ev_done()
            elif abs(tz - tk) <= 1e-9:
                pass  # measurement exactly at the IMU time: handled after the closing propagate below
            if kind == 0:
                vis, xt = pl
                if abs(tz - tk) <= 1e-9 and not at_tk:
                    tb.propagate(uk, tk, sigma=imu_sigma, sample=k)
                    info["n_prop"] += 1
                    at_tk = True
                    ev_done()
                ids = [v[0] for v in vis]
                dropped = [i for i in tracked if i not in ids]
                tb.keep_only_features(ids)
                info["n_keep"] += 1
                info["n_drop"] += len(dropped)
                tracked = [i for i in tracked if i in ids]
                ev_done()
                new = [v for v in vis if v[0] not in tracked]
                old = [v for v in vis if v[0] in tracked]
                for (fid, pix, depth) in new:  # initialised on arrival (vi_ekf_meas.cpp:140-147)
                    tb.update(abi.FEAT, pix, Rf, id=fid, depth=depth if depth_known else float("nan"), sigma=[pix_sigma, pix_sigma])
                    if len(tracked) < N:
                        tracked.append(fid)
                        info["n_init"] += 1
                    ev_done()
                for (fid, pix, depth) in reversed(old):  # LIFO tie order of the queue
                    tb.update(abi.FEAT, pix, Rf, id=fid, sigma=[pix_sigma, pix_sigma])
                    info["n_feat"] += 1
                    ev_done()
            else:
                if abs(tz - tk) <= 1e-9 and not at_tk:
                    tb.propagate(uk, tk, sigma=imu_sigma, sample=k)
                    info["n_prop"] += 1
                    at_tk = True
                    ev_done()
                tb.update(abi.ALT, [pl], [[alt_R]], sigma=[alt_sigma])
                info["n_alt"] += 1
                ev_done()
        if not at_tk:  # (a second propagate to the same time would return at once, vi_ekf.cpp:282: not emitted)
            tb.propagate(uk, tk, sigma=imu_sigma, sample=k)
            info["n_prop"] += 1
            ev_done()
        info["n_imu"] += 1
        if use_acc:
            # drag accelerometer update with the body-frame xy acceleration of this very IMU sample
            # (src/vi_ekf_ros.cpp:180-184): z is taken from the filter's last propagated input
            tb.update(abi.ACC, [0.0, 0.0], np.eye(2) * acc_R, from_imu=True)
            info["n_acc"] += 1
            ev_done()
        if stat_every and k % stat_every == 0:
            tb.stat(bucket, truth.x_true[k])
            bucket += 1
            info["n_stat"] += 1
            ev_done()
    info["live"] = np.array(live, dtype=np.int32)
    info["n_buckets"] = bucket
    return tb.events(), tb.id_pool(), info


# ---- algorithmic work model ------------------------------------------------------------
def flops_propagate(nfeat):
    return 427.5 * nfeat * nfeat + 3706.5 * nfeat + 6015.0


def flops_update(n, r):
    return (6 * r + 2) * n * (n + 1) / 2.0 + 8.0 * r * n + 30.0 * r * r


def tape_flops(events, live, first=0, count=None):
    """Algorithmic FP64 flops one filter spends on events[first:first+count] (every update counted as applied):
    returns (propagate flops, update flops, number of active updates)."""
    count = len(events) - first if count is None else count
    tot = 0.0
    totu = 0.0
    nupd = 0
    started = first > 0
    t_filter = None  # filter time: a propagate with |dt| < 1e-6 returns at once (vi_ekf.cpp:282) and costs nothing
    for i in range(first - 1, -1, -1):
        if events[i]["kind"] == abi.EV_PROP:
            t_filter = float(events[i]["t"])
            break
    for i in range(first, first + count):
        e = events[i]
        nf = int(live[i])
        if e["kind"] == abi.EV_PROP:
            t = float(e["t"])
            if started and (t_filter is None or abs(t - t_filter) >= 1e-6):
                tot += flops_propagate(nf)
            started = True
            t_filter = t
        elif e["kind"] == abi.EV_UPD:
            if e["type"] == abi.FEAT and i > 0 and live[i] != live[i - 1]:
                continue  # initialisation of a new feature, no update
            totu += flops_update(16 + 3 * nf, int(e["rdim"]))
            nupd += 1
    return tot, totu, nupd


def hbm_bytes_per_launch(nfeat):
    """x and packed-symmetric P of one filter, read once and written once by a fused launch."""
    n, m = 16 + 3 * nfeat, 17 + 5 * nfeat
    return 2 * (8 * m + 8 * n * (n + 1) // 2) + 48


# ---- the benchmark workload (BASELINE.json configs[1..4]): one builder for bench.py and for the parity tests -----------
BENCH_SEED = 0x5EED
BENCH_BIAS = [0.5, 0.5, 0.5, 0.05, 0.05, 0.05]  # sim_params.yaml:150-155 accel_init_stdev / gyro_init_stdev
BENCH_IMU_PER_STEP = 20


def bench_workload(N, n_slices, imu_per_step=BENCH_IMU_PER_STEP):
    """The tape bench.py times: the waypoint flight at 200 Hz IMU / 30 Hz camera with up to N tracked features
    (keep_only_features churn every frame), a drag-accelerometer update every IMU sample, the altimeter at 25 Hz, per-instance
    Philox IMU / pixel / altimeter noise and a constant per-instance IMU bias, one STAT event per slice of `imu_per_step`
    IMU samples.  use_keyframe_reset = 0: the altimeter measures the global altitude, which the reference's keyframe reset
    is inconsistent with (see build_tape).  Run it with seed = BENCH_SEED, bias_range = BENCH_BIAS, perturb_x0(BENCH_SEED + 1).
    Returns (cfg, events, id_pool, info, [(first, end)] event range of each slice)."""
    cfg = abi.default_config()
    cfg.use_keyframe_reset = 0
    truth = simulate(cfg, N, n_imu=n_slices * imu_per_step)
    ev, idp, info = build_tape(truth, cfg, N, stat_every=imu_per_step, depth_known=True)
    ends = [i + 1 for i in range(len(ev)) if ev[i]["kind"] == abi.EV_STAT]
    starts = [0] + ends[:-1]
    return cfg, ev, idp, info, list(zip(starts, ends))
