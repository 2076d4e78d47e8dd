"""TEST INFRASTRUCTURE -- second, independent restatement of the reference filter core in numpy float64.

SURVEY.md 8c asks for a ladder of checkers because the reference holds no x / P golden vector and cannot be compiled in this
image: (1) the C++ oracle (oracle/viekf_oracle.cpp, dense loops, function by function), (2) THIS file -- written separately
from the formulas of the reference's sources, vectorised with numpy (dense A, G, H matrices, `@` products, numpy's LAPACK
inverse), sharing no code with (1) -- and (3) symbolic Jacobians (tests/test_np_oracle.py, sympy).  Two restatements that
were written independently and agree to 1e-12 over 1000 steps shrink the room for a transcription error in either.

Every function cites the reference lines it follows.  Only tests/ may import this module.
"""
import numpy as np

GRAVITY = np.array([0.0, 0.0, 9.80665])  # include/vi_ekf.h:70-74
xPOS, xVEL, xATT, xB_A, xB_G, xMU, xZ = 0, 3, 6, 10, 13, 16, 17
dxPOS, dxVEL, dxATT, dxB_A, dxB_G, dxMU, dxZ = 0, 3, 6, 9, 12, 15, 16
ACC, ALT, ATT, POS, VEL, QZETA, FEAT, PIXEL_VEL, DEPTH, INV_DEPTH = range(10)
MEAS_SUCCESS, MEAS_GATED, MEAS_NAN, MEAS_INVALID, MEAS_NEW_FEATURE = range(5)
E_X, E_Y, E_Z = np.eye(3)


# ---- lib/geometry/include/geometry/quat.h ---------------------------------------------------------------------------
def skew(v):
    return np.array([[0.0, -v[2], v[1]], [v[2], 0.0, -v[0]], [-v[1], v[0], 0.0]])


def q_R(q):  # passive rotation matrix, quat.h:306-322
    w, x, y, z = q
    return np.array([[1 - 2 * y * y - 2 * z * z, 2 * x * y + 2 * w * z, 2 * x * z - 2 * w * y],
                     [2 * x * y - 2 * w * z, 1 - 2 * x * x - 2 * z * z, 2 * y * z + 2 * w * x],
                     [2 * x * z + 2 * w * y, 2 * y * z - 2 * w * x, 1 - 2 * x * x - 2 * y * y]])


def q_rotp(q, v):  # R v, quat.h:362-375
    return q_R(q) @ v


def q_rota(q, v):  # R^T v, quat.h:377-391
    return q_R(q).T @ v


def q_otimes(a, b):  # quat.h:419-428
    aw, av = a[0], a[1:]
    bw, bv = b[0], b[1:]
    return np.concatenate([[aw * bw - av @ bv], aw * bv + bw * av + np.cross(av, bv)])


def q_exp(v):  # quat.h:128-144
    n = np.linalg.norm(v)
    if n > 1e-4:
        return np.concatenate([[np.cos(n / 2)], np.sin(n / 2) * v / n])
    q = np.concatenate([[1.0], v / 2])
    return q / np.linalg.norm(q)


def q_log(q):  # quat.h:146-162
    n = np.linalg.norm(q[1:])
    if n < 1e-8:
        return np.zeros(3)
    return 2.0 * np.arctan2(n, q[0]) * q[1:] / n


def q_inv(q):
    return np.array([q[0], -q[1], -q[2], -q[3]])


def q_boxplus(q, d):  # quat.h:430-434
    return q_otimes(q, q_exp(d))


def q_boxminus(a, b):  # quat.h:436-445
    d = q_otimes(q_inv(b), a)
    if d[0] < 0:
        d = -d
    return q_log(d)


def q_from_two_unit_vectors(u, v):  # quat.h:233-255
    d = u @ v
    if -0.99999999 < d < 0.99999999:
        invs = 1.0 / np.sqrt(2.0 * (1.0 + d))
        q = np.concatenate([[0.5 / invs], np.cross(u, v * invs)])
        return q / np.linalg.norm(q)
    if d < -0.99999999:
        return np.array([0.0, 1.0, 0.0, 0.0])
    return np.array([1.0, 0.0, 0.0, 0.0])


def q_from_euler(roll, pitch, yaw):  # quat.h:196-212
    cp, ct, cs = np.cos(roll / 2), np.cos(pitch / 2), np.cos(yaw / 2)
    sp, st, ss = np.sin(roll / 2), np.sin(pitch / 2), np.sin(yaw / 2)
    return np.array([cp * ct * cs + sp * st * ss, sp * ct * cs - cp * st * ss, cp * st * cs + sp * ct * ss, cp * ct * ss - sp * st * cs])


def q_roll(q):
    return np.arctan2(2.0 * (q[0] * q[1] + q[2] * q[3]), 1.0 - 2.0 * (q[1] ** 2 + q[2] ** 2))


def q_pitch(q):
    v = 2.0 * (q[0] * q[2] - q[1] * q[3])
    return np.copysign(1.0, v) * np.pi / 2.0 if abs(v) > 1.0 else np.arcsin(v)


def q_yaw(q):
    return np.arctan2(2.0 * (q[0] * q[3] + q[1] * q[2]), 1.0 - 2.0 * (q[2] ** 2 + q[3] ** 2))


# ---- include/math_helper.h:14-48 ------------------------------------------------------------------------------------
def zeta_of(q):
    return q_rota(q, E_Z)


def T_zeta(q):
    return np.column_stack([q_rota(q, E_X), q_rota(q, E_Y)])


def qz_boxplus(q, d2):
    return q_otimes(q_exp(T_zeta(q) @ d2), q)


def qz_boxminus(qj, qi):
    zi, zj = zeta_of(qi), zeta_of(qj)
    if np.linalg.norm(zi - zj) > 1e-8:
        s = np.cross(zi, zj)
        s = s / np.linalg.norm(s)
        return T_zeta(qi).T @ (np.arccos(zi @ zj) * s)
    return np.zeros(2)


class NpVIEKF:
    """vi_ekf::VIEKF (include/vi_ekf.h:243-337) with NUM_FEATURES = NF, in-order calls only (no rewind queue)."""

    def __init__(self, cfg, NF):
        # VIEKF::load / init, src/vi_ekf/vi_ekf.cpp:64-155
        self.NF, self.n_max, self.m_max = NF, 16 + 3 * NF, 17 + 5 * NF
        self.x = np.zeros(self.m_max)
        self.x[:17] = np.array(cfg.x0[:])
        self.P = np.zeros((self.n_max, self.n_max))
        self.P[:16, :16] = np.diag(cfg.P0[:])
        self.Qx = np.zeros((self.n_max, self.n_max))
        self.Qx[:16, :16] = np.diag(cfg.Qx[:])
        lam = np.zeros(self.n_max)
        lam[:16] = cfg.lambda_[:]
        self.P0_feat = np.diag(cfg.P0_feat[:])
        for i in range(NF):
            sl = slice(16 + 3 * i, 19 + 3 * i)
            self.P[sl, sl] = self.P0_feat
            self.Qx[sl, sl] = np.diag(cfg.Qx_feat[:])
            lam[sl] = cfg.lambda_feat[:]
        self.Qu = np.diag(cfg.Qu[:])
        one = np.ones(self.n_max)
        self.Lambda = np.outer(one, lam) + np.outer(lam, one) - np.outer(lam, lam)  # vi_ekf.cpp:83,146
        self.lam = lam
        self.cam_center = np.array(cfg.cam_center[:])
        self.cam_F = np.array([[cfg.focal_len[0], 0, 0], [0, cfg.focal_len[1], 0]])
        self.q_b_c, self.p_b_c, self.q_b_u = np.array(cfg.q_b_c[:]), np.array(cfg.p_b_c[:]), np.array(cfg.q_b_u[:])
        self.min_depth = cfg.min_depth
        self.use_drag, self.use_partial, self.use_kfr = bool(cfg.use_drag_term), bool(cfg.use_partial_update), bool(cfg.use_keyframe_reset)
        self.kf_thresh = cfg.keyframe_overlap_threshold
        self.t, self.started = float("nan"), False
        self.len = 0
        self.next_id = 0
        self.ids = []
        self.kf_ids = []
        self.node_t, self.node_q = np.zeros(3), np.array([1.0, 0, 0, 0])
        self.ub_last = np.zeros(6)
        self.n_gated = 0

    # ---- vi_ekf_helper.cpp:88-111 ----
    def boxplus(self, x, dx):
        out = x.copy()
        out[0:6] = x[0:6] + dx[0:6]
        out[6:10] = q_boxplus(x[6:10], dx[6:9])
        out[10:17] = x[10:17] + dx[9:16]
        for i in range(self.len):
            xs, ds = 17 + 5 * i, 16 + 3 * i
            out[xs:xs + 4] = qz_boxplus(x[xs:xs + 4], dx[ds:ds + 2])
            out[xs + 4] = x[xs + 4] + dx[ds + 2]
        return out

    def fix_depth(self):  # vi_ekf_helper.cpp:128-156
        reset = 1.0 / (2.0 * self.min_depth)
        for i in range(self.len):
            xr, dr = 17 + 5 * i + 4, 16 + 3 * i + 2
            if self.x[xr] != self.x[xr]:
                self.x[xr] = reset
            if self.x[xr] < 0.0:
                err = reset - self.x[xr]
                self.P[dr, dr] += err * err
                self.x[xr] = reset
            elif self.x[xr] > 1e2:
                self.P[dr, dr] = self.P0_feat[2, 2]
                self.x[xr] = reset

    # ---- vi_ekf_dyn.cpp:14-135 ----
    def dynamics(self, x, ub):
        n = self.n_max
        dx, A, G = np.zeros(n), np.zeros((n, n)), np.zeros((n, 6))
        vel, q = x[3:6], x[6:10]
        acc, omega, mu = ub[0:3] - x[10:13], ub[3:6] - x[13:16], x[16]
        R = q_R(q)
        gB = R @ GRAVITY
        vxy = np.array([vel[0], vel[1], 0.0])
        dx[0:3] = R.T @ vel
        if self.use_drag:
            dx[3:6] = np.array([0, 0, acc[2]]) + gB - np.cross(omega, vel) - mu * vxy
        else:
            dx[3:6] = acc + gB - np.cross(omega, vel)
        dx[6:9] = omega
        A[0:3, 3:6] = R.T
        A[0:3, 6:9] = -R.T @ skew(vel)
        A[3:6, 3:6] = -skew(omega)
        A[3:6, 6:9] = skew(gB)
        A[3:6, 12:15] = -skew(vel)
        A[6:9, 6:9] = -skew(omega)
        A[6:9, 12:15] = -np.eye(3)
        G[3:6, 3:6] = -skew(vel)
        G[6:9, 3:6] = -np.eye(3)
        if self.use_drag:
            A[3:6, 3:6] += -mu * np.diag([1.0, 1.0, 0.0])
            A[3:6, 9:12] = -np.diag([0.0, 0.0, 1.0])
            A[3:6, 15] = -vxy
            G[3:6, 0:3] = -np.diag([0.0, 0.0, 1.0])
        else:
            A[3:6, 9:12] = -np.eye(3)
            G[3:6, 0:3] = -np.eye(3)
        Rbc = q_R(self.q_b_c)
        vc = Rbc @ (vel + np.cross(omega, self.p_b_c))
        oc = Rbc @ omega
        Sp = skew(self.p_b_c)
        for i in range(self.len):
            xs, ds = 17 + 5 * i, 16 + 3 * i
            qz, rho = x[xs:xs + 4], x[xs + 4]
            ze, T = zeta_of(qz), T_zeta(qz)
            Sz, Sv = skew(ze), skew(vc)
            w = oc + rho * np.cross(ze, vc)
            dx[ds:ds + 2] = -T.T @ w
            dx[ds + 2] = rho * rho * (ze @ vc)
            A[ds:ds + 2, 3:6] = -rho * T.T @ Sz @ Rbc
            A[ds:ds + 2, 12:15] = -T.T @ (rho * Sz @ Rbc @ Sp - Rbc)
            A[ds:ds + 2, ds:ds + 2] = -T.T @ (skew(w) + rho * Sv @ Sz) @ T
            A[ds:ds + 2, ds + 2] = -T.T @ np.cross(ze, vc)
            A[ds + 2, 3:6] = rho * rho * ze @ Rbc
            A[ds + 2, 12:15] = rho * rho * ze @ Rbc @ Sp
            A[ds + 2, ds:ds + 2] = rho * rho * ze @ Sv @ T
            A[ds + 2, ds + 2] = 2 * rho * (ze @ vc)
            G[ds:ds + 3, 3:6] = A[ds:ds + 3, 12:15]
        return dx, A, G

    # ---- vi_ekf.cpp:262-318 ----
    def propagate_state(self, u, t):
        ub = np.concatenate([q_rota(self.q_b_u, u[0:3]), q_rota(self.q_b_u, u[3:6])])
        self.ub_last = ub
        if not self.started:
            self.started, self.t = True, t
            return
        dt = t - self.t
        if abs(dt) < 1e-6 or dt < 0:
            return
        dx, A, G = self.dynamics(self.x, ub)
        eye = np.eye(self.n_max)
        self.x = self.boxplus(self.x, dx * dt)
        AA = A @ A
        Gd = (eye + A * dt / 2.0 + AA * dt * dt / 6.0) @ G * dt
        Ad = eye + A * dt + AA * dt * dt / 2.0
        self.P = Ad @ self.P @ Ad.T + Gd @ self.Qu @ Gd.T + self.Qx
        self.t = t
        self.fix_depth()

    # ---- vi_ekf_meas.cpp:281-395 ----
    def h(self, mtype, x, fid):
        H = np.zeros((3, self.n_max))
        if mtype == ACC:
            if self.use_drag:
                mu, vel = x[16], x[3:6]
                zh = np.array([-mu * vel[0] + x[10], -mu * vel[1] + x[11]])
                H[0, 3], H[1, 4] = -mu, -mu
                H[0, 9], H[1, 10] = 1.0, 1.0
                H[0, 15], H[1, 15] = -vel[0], -vel[1]
                return zh, H[:2]
            gB = q_rotp(x[6:10], GRAVITY)
            H[0:3, 6:9] = skew(-gB)
            H[0:3, 9:12] = np.eye(3)
            return x[10:13] - gB, H
        if mtype == ALT:
            H[0, 2] = -1.0
            return np.array([-x[2]]), H[:1]
        if mtype == ATT:
            H[0:3, 6:9] = np.eye(3)
            return x[6:10].copy(), H
        if mtype == POS:
            H[0:3, 0:3] = np.eye(3)  # (the reference indexes H with xPOS = dxPOS = 0)
            return x[0:3].copy(), H
        if mtype == VEL:
            H[0:3, 3:6] = np.eye(3)
            return x[3:6].copy(), H
        i = self.ids.index(fid)
        xs, ds = 17 + 5 * i, 16 + 3 * i
        if mtype == QZETA:
            H[0:2, ds:ds + 2] = np.eye(2)
            return x[xs:xs + 4].copy(), H[:2]
        if mtype == FEAT:
            qz = x[xs:xs + 4]
            ze, T = zeta_of(qz), T_zeta(qz)
            ez = ze[2]
            zh = self.cam_F @ ze / ez + self.cam_center
            H[0:2, ds:ds + 2] = (1.0 / ez) * self.cam_F @ (np.outer(ze, E_Z) / ez - np.eye(3)) @ skew(ze) @ T
            return zh, H[:2]
        if mtype == DEPTH:
            rho = x[xs + 4]
            H[0, ds + 2] = -1.0 / (rho * rho)
            return np.array([1.0 / rho]), H[:1]
        if mtype == INV_DEPTH:
            H[0, ds + 2] = 1.0
            return np.array([x[xs + 4]]), H[:1]
        return np.zeros(2), H[:2]  # PIXEL_VEL: empty stub (vi_ekf_meas.cpp:388-395)

    # ---- add_measurement bookkeeping + update, vi_ekf_meas.cpp:130-147,196-278 (in-order tape semantics) ----
    def update(self, mtype, z, R, active=True, fid=-1, depth=float("nan"), replay=False):
        z = np.atleast_1d(np.asarray(z, dtype=float))
        R = np.atleast_2d(np.asarray(R, dtype=float))
        if np.isnan(z).any():
            return MEAS_NAN
        if mtype in (FEAT, QZETA, DEPTH, INV_DEPTH) and fid not in self.ids:
            if mtype == FEAT and fid >= 0 and not replay:
                self.init_feature(z[:2], fid, depth)
                return MEAS_NEW_FEATURE
            return MEAS_INVALID
        zh, H = self.h(mtype, self.x, fid)
        if mtype == QZETA:
            res = qz_boxminus(z, zh)
        elif mtype == ATT:
            res = q_boxminus(z, zh)
        else:
            res = z[:len(zh)] - zh
        r = R.shape[0]
        H = H[:r]
        if active:
            S = H @ self.P @ H.T + R
            Si = np.linalg.inv(S)
            if res @ Si @ res > 9.0:
                self.n_gated += 1
                return MEAS_GATED
            K = self.P @ H.T @ Si
            if not (np.isnan(K).any() or np.isnan(H).any()):
                corr = K @ res
                IKH = np.eye(self.n_max) - K @ H
                Pn = IKH @ self.P @ IKH.T + K @ R @ K.T
                if self.use_partial:
                    self.x = self.boxplus(self.x, self.lam * corr)
                    self.P = self.P + self.Lambda * (Pn - self.P)
                else:
                    self.x = self.boxplus(self.x, corr)
                    self.P = Pn
        self.fix_depth()
        return MEAS_SUCCESS

    # ---- vi_ekf_feat.cpp:6-142 ----
    def init_feature(self, l, fid, depth=float("nan")):
        if self.len >= self.NF:
            return False
        lc = np.asarray(l, dtype=float) - self.cam_center
        ze = np.array([lc[0], lc[1] * (self.cam_F[1, 1] / self.cam_F[0, 0]), self.cam_F[0, 0]])
        ze = ze / np.linalg.norm(ze)
        qz = q_from_two_unit_vectors(E_Z, ze)
        d0 = 2.0 * self.min_depth if depth != depth else depth
        self.ids.append(self.next_id)  # the caller's id is ignored
        self.next_id += 1
        self.len += 1
        xs, ds = 17 + 5 * (self.len - 1), 16 + 3 * (self.len - 1)
        self.x[xs:xs + 4], self.x[xs + 4] = qz, 1.0 / d0
        self.P[ds:ds + 3, :ds + 3] = 0.0
        self.P[:ds + 3, ds:ds + 3] = 0.0
        self.P[ds:ds + 3, ds:ds + 3] = self.P0_feat
        return True

    def clear_feature(self, fid):
        i = self.ids.index(fid)
        self.ids.pop(i)
        self.len -= 1
        keep_x = np.r_[0:17 + 5 * i, 17 + 5 * (i + 1):self.m_max]
        keep_d = np.r_[0:16 + 3 * i, 16 + 3 * (i + 1):self.n_max]
        x = np.zeros(self.m_max)
        x[:len(keep_x)] = self.x[keep_x]
        P = np.zeros((self.n_max, self.n_max))
        P[:len(keep_d), :len(keep_d)] = self.P[np.ix_(keep_d, keep_d)]
        nl, ml = 16 + 3 * self.len, 17 + 5 * self.len
        x[ml:] = 0.0
        P[nl:, :] = 0.0
        P[:, nl:] = 0.0
        self.x, self.P = x, P

    def keep_only_features(self, features):
        features = [int(f) for f in features]
        overlap = sum(1 for f in self.ids if f in features and self.use_kfr and f in self.kf_ids)
        for f in [f for f in self.ids if f not in features]:
            self.clear_feature(f)
        if self.use_kfr and len(self.kf_ids) > 0 and overlap / len(self.kf_ids) < self.kf_thresh:
            self.keyframe_reset()
            self.kf_ids = list(features)
        elif self.use_kfr and len(self.kf_ids) == 0:
            self.kf_ids = list(features)

    # ---- vi_ekf_kfr.cpp:56-157 ----
    def keyframe_reset(self):
        et = self.x[0:3].copy()
        self.x[0:3] = 0.0
        q = self.x[6:10]
        yaw, roll, pitch = q_yaw(q), q_roll(q), q_pitch(q)
        qe = q_from_euler(0, 0, yaw)
        self.x[6:10] = q_from_euler(roll, pitch, 0.0)
        cp, sp, tt = np.cos(roll), np.sin(roll), np.tan(pitch)
        N = np.eye(self.n_max)
        N[0:3, 0:3] = 0.0
        N[6:9, 6:9] = np.array([[1, sp * tt, cp * tt], [0, cp * cp, -cp * sp], [0, -cp * sp, sp * sp]])
        self.P = N @ self.P @ N.T
        self.node_t = self.node_t + q_rota(self.node_q, et)  # xform.h:236-244: global = node * edge
        self.node_q = q_otimes(self.node_q, qe)

    # ---- in-order event tape (include/viekf_batch.h), no generated noise ----
    def run_tape(self, events, id_pool):
        res = np.full(len(events), 255, dtype=np.uint8)
        for k, e in enumerate(events):
            kind = int(e["kind"])
            if kind == 0:
                self.propagate_state(np.array(e["z"][:6]), float(e["t"]))
            elif kind == 1:
                zd, rd = int(e["zdim"]), int(e["rdim"])
                z = self.ub_last[:zd].copy() if int(e["aux"]) == 1 else np.array(e["z"][:zd])
                R = np.array(e["R"][:rd * rd]).reshape(rd, rd).T
                res[k] = self.update(int(e["type"]), z, R, bool(e["active"]), int(e["id"]), float(e["z"][4]), replay=int(e["aux"]) == 2)
            elif kind == 2:
                res[k] = 1 if self.init_feature(np.array(e["z"][:2]), int(e["id"]), float(e["z"][4])) else 0
            elif kind == 3:
                if int(e["id"]) in self.ids:
                    self.clear_feature(int(e["id"]))
            elif kind == 4:
                self.keep_only_features(id_pool[int(e["ids_off"]):int(e["ids_off"]) + int(e["n_ids"])])
            elif kind == 6:
                self.keyframe_reset()
        return res
