"""The reference's simulated vehicle and waypoint controller, restated.

lib/multirotor_sim flies its multirotor with a nonlinear velocity controller (include/multirotor_sim/nlc.h:54-132), three
PID loops on roll / pitch / yaw rate (include/multirotor_sim/pid.h:40-96), a hover-throttle observer and a waypoint manager
(src/controller.cpp:18-62,229-279) on top of force / torque rigid-body dynamics integrated by RK4 (src/dynamics.cpp:50-106)
and an inner angular-rate loop (src/simulator.cpp:807-819).  This module restates exactly that chain with the gains of
params/sim_params.yaml, so that the synthetic IMU / camera streams of `sim.simulate(controller="reference")` come from the
same closed loop as the reference's.  Known answer held by the reference (src/test/test_reference_controller.cpp:79-83):
with `seed: 1`, dt = 0.01 and the six waypoints of that test the waypoint index changes at 0.01 / 5.25 / 16.45 / 30.25 / 44.32 s
(+- 0.1 s) -- tests/test_refsim.py.

State conventions (include/multirotor_sim/state.h): p inertial NED, q inertial -> body ([w, x, y, z], rotp = R v passive),
v and w in the body frame.  Process noise uses numpy's generator, not std::default_random_engine, so noisy runs agree with the
reference statistically, not sample for sample.
"""
import numpy as np

G = 9.80665

# params/sim_params.yaml (reference lib/multirotor_sim/params/sim_params.yaml:52-104)
DEFAULT_PARAMS = dict(
    throttle_eq=0.5, mass=1.0, max_thrust=19.6133, max_torque=(0.30625, 0.30625, 0.1), kp_w=(1.0, 1.0, 1.0), kd_w=(0.0, 0.0, 0.0),
    waypoint_threshold=0.1, waypoint_velocity_threshold=0.5, drag_constant=0.1, angular_drag_constant=0.01,
    Kp=(1.0, 1.0, 1.0), Kd=(0.0, 0.0, 0.0), Kv=(2.0, 2.0, 2.0), sh_kv=50.0, sh_ks=0.1,
    roll_kp=10.0, roll_ki=0.0, roll_kd=1.0, pitch_kp=10.0, pitch_ki=0.0, pitch_kd=1.0, yaw_rate_kp=1.0, yaw_rate_ki=0.0, yaw_rate_kd=0.0,
    max_roll=1.0, max_pitch=1.0, max_yaw_rate=1.0, max_throttle=1.0, max_vel=5.0, max_tau_x=1.0, max_tau_y=1.0, max_tau_z=1.0,
    inertia=(0.1, 0.1, 0.1), enable_dynamics_noise=True,
    dyn_noise=(0, 0, 0, 0, 0, 0, 0.02, 0.02, 0.02, 0.01, 0.01, 0.01),  # POS, ATT, VEL, OMEGA
)
# the waypoints of sim_params.yaml:14-21 (also the ones test_reference_controller.cpp sets)
REFERENCE_WAYPOINTS = np.array([[10, 0, -5, .705], [10, 10, -4, 3.0], [-10, 10, -5.5, -1.5], [-10, -7, -5, .80], [-8, -12, -4.5, -2.3],
                                [8, -5, -4, 1.7]], dtype=float)


# ---- quaternion helpers, conventions of lib/geometry quat.h ---------------------------------------------------------
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


def q_from_euler(r, p, y):
    cp, ct, cs = np.cos(r / 2), np.cos(p / 2), np.cos(y / 2)
    sp, st, ss = np.sin(r / 2), np.sin(p / 2), np.sin(y / 2)
    return np.array([cp * ct * cs + sp * st * ss, sp * ct * cs - cp * st * ss, cp * st * cs + sp * ct * ss, cp * ct * ss - sp * st * cs])


def q_roll(q):
    return np.arctan2(2.0 * (q[0] * q[1] + q[2] * q[3]), 1.0 - 2.0 * (q[1] * q[1] + q[2] * q[2]))


def q_pitch(q):
    v = 2.0 * (q[0] * q[2] - q[1] * q[3])
    if abs(v) > 1.0:
        return np.copysign(1.0, v) * np.pi / 2.0
    return np.arcsin(v)


def q_yaw(q):
    return np.arctan2(2.0 * (q[0] * q[3] + q[1] * q[2]), 1.0 - 2.0 * (q[2] * q[2] + q[3] * q[3]))


def sat(x, hi, lo):
    return hi if x > hi else lo if x < lo else x


def R_v_to_b(phi, theta, psi):
    """3-2-1 Euler rotation from the vehicle frame to the body frame (include/multirotor_sim/utils.h:163-199)."""
    c, s = np.cos, np.sin
    R1 = np.array([[c(psi), s(psi), 0], [-s(psi), c(psi), 0], [0, 0, 1.0]])
    R2 = np.array([[c(theta), 0, -s(theta)], [0, 1.0, 0], [s(theta), 0, c(theta)]])
    R3 = np.array([[1.0, 0, 0], [0, c(phi), s(phi)], [0, -s(phi), c(phi)]])
    return R3 @ R2 @ R1


class State:
    """p, q, v, w (include/multirotor_sim/state.h)."""

    def __init__(self, p=(0, 0, 0), q=(1, 0, 0, 0), v=(0, 0, 0), w=(0, 0, 0)):
        self.p, self.q, self.v, self.w = (np.array(a, dtype=float) for a in (p, q, v, w))

    def copy(self):
        return State(self.p, self.q, self.v, self.w)

    def plus(self, d, s=1.0):
        """x [+] d s with d = (dp, dq, dv, dw)."""
        return State(self.p + d[0] * s, q_mul(self.q, q_exp(d[1] * s)), self.v + d[2] * s, self.w + d[3] * s)


class PID:
    """include/multirotor_sim/pid.h:40-96."""

    def __init__(self, kp, ki, kd, mx, tau=0.05):
        self.kp, self.ki, self.kd, self.max, self.tau = kp, ki, kd, mx, tau
        self.integrator = self.differentiator = self.prev_x = 0.0

    def run(self, dt, x, x_c, update_integrator, xdot=None):
        if xdot is None:  # dirty derivative (pid.h:40-58)
            if dt > 0.0001:
                self.differentiator = (2.0 * self.tau - dt) / (2.0 * self.tau + dt) * self.differentiator + 2.0 / (2.0 * self.tau + dt) * (x - self.prev_x)
                xdot = self.differentiator
            else:
                xdot = 0.0
            self.prev_x = x
        error = x_c - x
        p_term = error * self.kp
        d_term = self.kd * xdot if self.kd > 0.0 else 0.0
        i_term = 0.0
        if self.ki > 0.0 and update_integrator:
            self.integrator += error * dt
            i_term = self.ki * self.integrator
        u = p_term - d_term + i_term
        u_sat = self.max if u > self.max else -self.max if u < -self.max else u
        if u != u_sat and abs(i_term) > abs(u - p_term + d_term) and self.ki > 0.0:
            self.integrator = (u_sat - p_term + d_term) / self.ki
        return u_sat


class ReferenceController:
    """Waypoint manager + nonlinear velocity controller + roll / pitch / yaw-rate PID + hover-throttle observer
    (src/controller.cpp:18-62,229-279, include/multirotor_sim/nlc.h:54-132)."""

    def __init__(self, waypoints=REFERENCE_WAYPOINTS, params=None):
        P = dict(DEFAULT_PARAMS)
        P.update(params or {})
        self.P = P
        self.waypoints = np.asarray(waypoints, dtype=float)
        self.Kp, self.Kd, self.Kv = (np.diag(P[k]) for k in ("Kp", "Kd", "Kv"))
        self.sh = P["throttle_eq"]
        self.sh_inv_hat = 1.0 / self.sh
        self.roll = PID(P["roll_kp"], P["roll_ki"], P["roll_kd"], P["max_tau_x"])
        self.pitch = PID(P["pitch_kp"], P["pitch_ki"], P["pitch_kd"], P["max_tau_y"])
        self.yaw_rate = PID(P["yaw_rate_kp"], P["yaw_rate_ki"], P["yaw_rate_kd"], P["max_tau_z"])
        self.xhat, self.xc = State(), State()
        self.vhat = np.zeros(3)
        self.dhat = np.zeros(3)
        self.s_prev = 0.0
        self.prev_time = 0.0
        self.initialized = False
        self.current_waypoint_id = -1

    def _set_waypoint(self, i):
        wpt = self.waypoints[i]
        self.xc.p = wpt[:3].copy()
        self.xc.q = q_from_euler(q_roll(self.xc.q), q_pitch(self.xc.q), wpt[3])
        self.current_waypoint_id = i

    def get_commanded_state(self, t):
        """updateWaypointManager (controller.cpp:229-270): next waypoint once position + heading error < threshold and slow enough."""
        if not self.initialized:
            self.initialized = True
            self._set_waypoint(0)
        wpt = self.waypoints[self.current_waypoint_id]
        err = np.array([wpt[0] - self.xhat.p[0], wpt[1] - self.xhat.p[1], wpt[2] - self.xhat.p[2], wpt[3] - q_yaw(self.xhat.q)])
        if err[3] > np.pi:
            err[3] -= 2.0 * np.pi
        elif err[3] < -np.pi:
            err[3] += 2.0 * np.pi
        if np.linalg.norm(err) < self.P["waypoint_threshold"] and np.linalg.norm(self.xhat.v) < self.P["waypoint_velocity_threshold"]:
            self._set_waypoint((self.current_waypoint_id + 1) % len(self.waypoints))
        return self.xc

    def _nlc(self, dt):
        """NLC::computeControl, path types < 3 (nlc.h:54-132).  Returns the throttle; updates xc.v, xc.w[2], xc.q."""
        P, xhat, xc = self.P, self.xhat, self.xc
        e3 = np.array([0.0, 0.0, 1.0])
        psi = q_yaw(xhat.q)
        R_v_to_v1 = np.array([[np.cos(psi), np.sin(psi), 0], [-np.sin(psi), np.cos(psi), 0], [0, 0, 1.0]])
        vc = R_v_to_v1 @ self.Kp @ (xc.p - xhat.p)
        vmag = np.linalg.norm(vc)
        if vmag > P["max_vel"]:
            vc = vc * (P["max_vel"] / vmag)
        xc.v = vc
        wz = q_yaw(xc.q) - psi
        if wz > np.pi:
            wz -= 2 * np.pi
        if wz < -np.pi:
            wz += 2 * np.pi
        xc.w[2] = sat(wz, P["max_yaw_rate"], -P["max_yaw_rate"])
        R_v1_to_b = R_v_to_b(q_roll(xhat.q), q_pitch(xhat.q), 0.0)
        vhat = R_v1_to_b.T @ xhat.v
        self.dhat = self.dhat - self.Kd @ (xhat.v - vhat) * dt
        k_tilde = self.sh * (e3 - (1.0 / G) * (self.Kv @ (xc.v - vhat) - self.dhat))
        throttle = sat(float(e3 @ R_v1_to_b @ k_tilde), P["max_throttle"], 0.001)
        kd = k_tilde / throttle
        kd = kd / np.linalg.norm(kd)
        kTkd = kd[2]
        tilt = np.arccos(kTkd) if abs(kTkd - 1.0) > 1.0e-6 else 0.0
        if tilt < 1e-6:
            qc = np.array([1.0, 0, 0, 0])
        else:
            kx = np.cross(e3, kd)
            qc = q_exp(tilt * kx / np.linalg.norm(kx))
        phi = sat(q_roll(qc), P["max_roll"], -P["max_roll"])
        theta = sat(q_pitch(qc), P["max_pitch"], -P["max_pitch"])
        xc.q = q_from_euler(phi, theta, q_yaw(xc.q))
        return throttle, R_v1_to_b

    def compute_control(self, t, x):
        """ReferenceController::computeControl (controller.cpp:18-62).  Returns u = [throttle 0..1, rate commands (3)]."""
        self.xhat = x.copy()
        dt = t - self.prev_time
        self.prev_time = t
        if dt < 1e-7:
            return np.zeros(4)
        throttle, R_v1_to_b = self._nlc(dt)
        xhat, xc = self.xhat, self.xc
        u = np.array([throttle,
                      self.roll.run(dt, q_roll(xhat.q), q_roll(xc.q), False, xhat.w[0]),
                      self.pitch.run(dt, q_pitch(xhat.q), q_pitch(xc.q), False, xhat.w[1]),
                      self.yaw_rate.run(dt, xhat.w[2], xc.w[2], False)])
        e3 = np.array([0.0, 0.0, 1.0])
        vhat_dot = G * (np.eye(3) - self.sh_inv_hat * self.s_prev * R_v1_to_b.T) @ e3 - np.cross(xhat.w, self.vhat) + self.P["sh_kv"] * (xhat.v - self.vhat)
        sh_inv_hat_dot = -self.P["sh_ks"] * G * self.s_prev * float((xhat.v - self.vhat) @ R_v1_to_b.T @ e3)
        self.vhat = self.vhat + vhat_dot * dt
        self.sh_inv_hat += sh_inv_hat_dot * dt
        self.sh = 1.0 / self.sh_inv_hat
        self.s_prev = throttle
        return u


class Vehicle:
    """Dynamics::f / run (src/dynamics.cpp:50-106) and Simulator::compute_low_level_control (src/simulator.cpp:807-819)."""

    def __init__(self, x0, params=None, seed=1):
        P = dict(DEFAULT_PARAMS)
        P.update(params or {})
        self.P = P
        self.x = x0.copy()
        self.J = np.diag(P["inertia"])
        self.Jinv = np.linalg.inv(self.J)
        self.Qsqrt = np.sqrt(np.array(P["dyn_noise"], dtype=float))
        self.rng = np.random.default_rng(seed)
        self.w_err_prev = np.zeros(3)
        self.kp_w, self.kd_w, self.max_torque = (np.array(P[k], dtype=float) for k in ("kp_w", "kd_w", "max_torque"))
        self.ft = np.zeros(4)

    def f(self, x, ft):
        P = self.P
        dp = q_rota(x.q, x.v)
        dv = -np.array([0.0, 0.0, 1.0]) * ft[0] / P["mass"] - P["drag_constant"] * x.v + q_rotp(x.q, np.array([0.0, 0.0, G])) - np.cross(x.w, x.v)
        dw = self.Jinv @ (ft[1:4] - np.cross(x.w, self.J @ x.w) - P["angular_drag_constant"] * x.w * x.w)
        return dp, x.w.copy(), dv, dw

    def imu(self, x, ft):
        """Specific force and angular rate in the body frame (dynamics.cpp:59-64 with p_b_u = 0, q_b_u applied by the caller)."""
        _, _, dv, _ = self.f(x, ft)
        return dv + np.cross(x.w, x.v) - q_rotp(x.q, np.array([0.0, 0.0, G])), x.w.copy()

    def low_level(self, u, dt):
        ft = np.zeros(4)
        ft[0] = sat(u[0] * self.P["max_thrust"], self.P["max_thrust"], 0.0)
        w_err = u[1:4] - self.x.w
        tau = self.kp_w * w_err - self.kd_w * (w_err - self.w_err_prev) / dt
        self.w_err_prev = w_err
        ft[1:4] = np.minimum(np.maximum(tau, -self.max_torque), self.max_torque)
        return ft

    def run(self, dt, ft):
        x = self.x
        self.ft = ft
        k1 = self.f(x, ft)
        k2 = self.f(x.plus(k1, dt / 2.0), ft)
        k3 = self.f(x.plus(k2, dt / 2.0), ft)
        k4 = self.f(x.plus(k3, dt), ft)
        d = [(a + 2.0 * b + 2.0 * c + e) * (dt / 6.0) for a, b, c, e in zip(k1, k2, k3, k4)]
        if self.P["enable_dynamics_noise"]:
            n = self.Qsqrt * self.rng.standard_normal(12) * dt
            d = [d[0] + n[0:3], d[1] + n[3:6], d[2] + n[6:9], d[3] + n[9:12]]
        self.x = x.plus(d)


def fly(tmax=60.0, dt=0.01, waypoints=REFERENCE_WAYPOINTS, x0=None, params=None, seed=1, on_step=None):
    """Simulator::run (src/simulator.cpp:100-121) without sensors: t += dt, commanded state, control, dynamics.
    Returns (times at which the waypoint index changed, final State).  `on_step(t, vehicle, controller)` is called after each step."""
    x0 = x0 or State(p=(0, 0, -5), q=(1, 0, 0, 0), v=(1, 0, 0), w=(0, 0, 0))
    veh = Vehicle(x0, params, seed)
    con = ReferenceController(waypoints, params)
    t, prev, times = 0.0, -1, []
    while t < tmax - dt / 2.0:
        t += dt
        con.get_commanded_state(t)
        u = con.compute_control(t, veh.x)
        veh.run(dt, veh.low_level(u, dt))
        if on_step is not None:
            on_step(t, veh, con)
        if con.current_waypoint_id != prev:
            prev = con.current_waypoint_id
            times.append(t)
    return times, veh.x
