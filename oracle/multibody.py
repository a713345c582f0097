"""multibody — the general Bullet multibody step for the exo_v3.urdf tree (minibullet `motor_model="full"`).

TEST INFRASTRUCTURE ONLY (oracle). Nothing in the product imports this file.

Restates, for a fixed-base tree of revolute/prismatic joints, what one
`btMultiBodyDynamicsWorld::stepSimulation` sub-step does when the only constraints are joint motors and
joint limits (no contacts): see the numbered list in oracle/minibullet.py's header. Reference call sites it stands
in for: Environment/Exoskeleton_env.py:113-118 (gravity, dt), :433 (stepSimulation);
Environment/Exoskeleton_sim_pybullet.py:18 (loadURDF, fixed base, no flags), :115-117 (POSITION_CONTROL, default gains).
Link masses / inertias / collision boxes: Simulation/exo_v3.urdf:3-13,22-33,42-53,62-73,82-93,102-113,121-132,...

Everything here is double precision and written for clarity, not speed (a step costs a few ms).
[bullet-memory]: Bullet's source is not available in this image; constants and ordering below are restated from its
published code and documented defaults. The purpose is to BOUND how far the product's closed-form tracker
(q+ = q + kp (target - q)) can be from the general algorithm, over every unverifiable choice (knobs).
"""
import math

import numpy as np


def _cross(a, b):
    return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]])


def bullet_quicksort_equal_keys(n):
    """Order in which btAlignedObjectArray::quickSort leaves n elements whose keys all compare equal.

    btMultiBodyDynamicsWorld sorts the world's multibody constraints by island id with this (unstable) quicksort
    before the solver creates rows from them; all 38 constraints of the one body share one island, so the comparator
    is always false and the partition step swaps symmetric pairs. Returns perm with sorted[k] = original[perm[k]]."""
    a = list(range(n))

    def qs(lo, hi):
        i, j = lo, hi
        # pivot value is never used by an always-false comparator
        while True:
            if i <= j:
                a[i], a[j] = a[j], a[i]
                i += 1
                j -= 1
            if not (i <= j):
                break
        if lo < j:
            qs(lo, j)
        if i < hi:
            qs(i, hi)

    if n > 1:
        qs(0, n - 1)
    return a


class TreeDynamics(object):
    def __init__(self, body, inertia_from="collision", linear_damping=0.04, angular_damping=0.04):
        assert inertia_from in ("collision", "file")
        self.b = body
        self.n = body.n
        self.k_lin, self.k_ang = linear_damping, angular_damping
        self.mass = np.array([ln.mass for ln in body.links])
        self.I_local = []
        for ln in body.links:
            diag = ln.inertia_file
            if inertia_from == "collision" and ln.boxes:
                # btCompoundShape::calculateLocalInertia: box inertia of the compound's AABB in the inertial frame
                lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
                for size, xyz, Rc in ln.boxes:
                    Rrel = ln.R_inertial.T @ Rc
                    c = ln.R_inertial.T @ (xyz - ln.com_xyz)
                    h = np.abs(Rrel) @ (0.5 * size)
                    lo, hi = np.minimum(lo, c - h), np.maximum(hi, c + h)
                l = hi - lo
                diag = ln.mass / 12.0 * np.array([l[1] ** 2 + l[2] ** 2, l[0] ** 2 + l[2] ** 2, l[0] ** 2 + l[1] ** 2])
            self.I_local.append(ln.R_inertial @ np.diag(diag) @ ln.R_inertial.T)
        self.anc = []            # ancestors-or-self of every link, root first
        for i, ln in enumerate(body.links):
            self.anc.append((self.anc[ln.parent] if ln.parent >= 0 else []) + [i])

    # -- kinematics -------------------------------------------------------------------------------------------------
    def kinematics(self, q):
        b = self.b
        R, p, ax, com, Iw = [None] * self.n, [None] * self.n, [None] * self.n, [None] * self.n, [None] * self.n
        base_R = np.eye(3)
        base_p = b.base_pos - base_R @ b.base_com
        for i, ln in enumerate(b.links):
            pR, pp = (base_R, base_p) if ln.parent < 0 else (R[ln.parent], p[ln.parent])
            Rj = pR @ ln.origin_R
            pj = pp + pR @ ln.origin_xyz
            a = ln.axis / np.linalg.norm(ln.axis)
            ax[i] = Rj @ a
            if ln.joint_type == "prismatic":
                R[i], p[i] = Rj, pj + ax[i] * q[i]
            else:
                c, s = math.cos(q[i]), math.sin(q[i])
                K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
                R[i], p[i] = Rj @ (np.eye(3) + s * K + (1 - c) * (K @ K)), pj
            com[i] = p[i] + R[i] @ ln.com_xyz
            Iw[i] = R[i] @ self.I_local[i] @ R[i].T
        return dict(R=R, p=p, ax=ax, com=com, Iw=Iw)

    def mass_matrix(self, kin):
        """M(q) = sum_i m_i Jv_i^T Jv_i + Jw_i^T I_i Jw_i (composite form; what ABA's factorisation inverts)."""
        n = self.n
        M = np.zeros((n, n))
        for i in range(n):
            Jv, Jw = np.zeros((3, n)), np.zeros((3, n))
            for j in self.anc[i]:
                if self.b.links[j].joint_type == "prismatic":
                    Jv[:, j] = kin["ax"][j]
                else:
                    Jw[:, j] = kin["ax"][j]
                    Jv[:, j] = _cross(kin["ax"][j], kin["com"][i] - kin["p"][j])
            M += self.mass[i] * (Jv.T @ Jv) + Jw.T @ kin["Iw"][i] @ Jw
        return M

    def bias(self, kin, qd, gravity):
        """Joint forces that hold qdd = 0: Coriolis/centrifugal + gyroscopic - gravity - damping (recursive Newton-Euler)."""
        n, b = self.n, self.b
        g = np.asarray(gravity, dtype=np.float64)
        w, al, v, a = [None] * n, [None] * n, [None] * n, [None] * n
        z = np.zeros(3)
        for i, ln in enumerate(b.links):
            pw, pal, pv, pa, pp = (z, z, z, z, None) if ln.parent < 0 else (w[ln.parent], al[ln.parent], v[ln.parent], a[ln.parent], kin["p"][ln.parent])
            r = z if pp is None else kin["p"][i] - pp
            v_i = pv + _cross(pw, r)
            a_i = pa + _cross(pal, r) + _cross(pw, _cross(pw, r))
            if ln.joint_type == "prismatic":
                w[i], al[i] = pw, pal
                v[i] = v_i + kin["ax"][i] * qd[i]
                a[i] = a_i + 2.0 * _cross(pw, kin["ax"][i]) * qd[i]
            else:
                w[i] = pw + kin["ax"][i] * qd[i]
                al[i] = pal + _cross(pw, kin["ax"][i]) * qd[i]
                v[i], a[i] = v_i, a_i
        f, m = [None] * n, [None] * n
        for i in range(n):
            rc = kin["com"][i] - kin["p"][i]
            vc = v[i] + _cross(w[i], rc)
            ac = a[i] + _cross(al[i], rc) + _cross(w[i], _cross(w[i], rc))
            Iw_ = kin["Iw"][i] @ w[i]
            # Bullet's per-link damping (btMultiBody linear/angular damping, both K1 and K2 terms)
            f_ext = self.mass[i] * g - self.mass[i] * vc * (self.k_lin + self.k_lin * np.linalg.norm(vc))
            t_ext = -Iw_ * (self.k_ang + self.k_ang * np.linalg.norm(w[i]))
            F = self.mass[i] * ac - f_ext
            N = kin["Iw"][i] @ al[i] + _cross(w[i], Iw_) - t_ext
            f[i] = F
            m[i] = N + _cross(rc, F)          # moment about the link-frame origin p_i
        tau = np.zeros(n)
        for i in range(n - 1, -1, -1):
            ln = b.links[i]
            tau[i] = kin["ax"][i] @ (f[i] if ln.joint_type == "prismatic" else m[i])
            if ln.parent >= 0:
                f[ln.parent] = f[ln.parent] + f[i]
                m[ln.parent] = m[ln.parent] + m[i] + _cross(kin["p"][i] - kin["p"][ln.parent], f[i])
        return tau


class FullStep(object):
    """One Bullet multibody sub-step with motor + limit rows (see module docstring)."""

    def __init__(self, body, dt, gravity, kp, kd, max_force=1.0e5, inertia_from="collision", iterations=50,
                 residual_threshold=1e-7, row_order="bullet", erp=0.2, limit_max_impulse=100.0,
                 default_motor_impulse=1.0, linear_damping=0.04, angular_damping=0.04):
        self.body, self.dt, self.gravity = body, dt, gravity
        self.kp, self.kd, self.max_force = kp, kd, max_force
        self.dyn = TreeDynamics(body, inertia_from, linear_damping, angular_damping)
        self.iterations, self.threshold, self.erp = iterations, residual_threshold, erp
        self.limit_max_impulse, self.default_motor_impulse = limit_max_impulse, default_motor_impulse
        n = body.n
        # world constraint list: joint-limit constraints in link order (URDF import), then createJointMotors' motors
        cons = [("limit", j) for j in range(n) if body.links[j].lower <= body.links[j].upper] + [("motor", j) for j in range(n)]
        if row_order == "bullet":
            cons = [cons[k] for k in bullet_quicksort_equal_keys(len(cons))]
        else:
            assert row_order == "natural"
        self.constraints = cons
        self.stats = dict(steps=0, sweeps=0, max_residual=0.0, cap_hits=0, limit_rows=0, max_prismatic=0.0)

    def step(self):
        b, dt, n = self.body, self.dt, self.body.n
        kin = self.dyn.kinematics(b.q)
        M = self.dyn.mass_matrix(kin)
        Minv = np.linalg.inv(M)
        qd = b.qd + dt * (Minv @ (-self.dyn.bias(kin, b.qd, self.gravity)))     # no applied joint torques on this path
        rows = []                      # (joint, sign, desired-velocity error, lower, upper)
        for kind, j in self.constraints:
            ln = b.links[j]
            if kind == "motor":
                if j in b.targets:     # POSITION_CONTROL (pybullet defaults: kp 0.1, kd 1, target velocity 0, maxForce 1e5)
                    desired = self.kp * (b.targets[j] - b.q[j]) / dt + qd[j] + self.kd * (0.0 - qd[j])
                    cap = self.max_force * dt
                else:                  # createJointMotors: velocity motor, kp 0, kd 1, target 0, max impulse 1
                    desired = qd[j] + 1.0 * (0.0 - qd[j])
                    cap = self.default_motor_impulse
                rows.append((j, 1.0, desired - qd[j], -cap, cap))
            else:
                for sign, pen in ((1.0, b.q[j] - ln.lower), (-1.0, ln.upper - b.q[j])):
                    if pen > 0:
                        continue
                    rows.append((j, sign, -pen * self.erp / dt - sign * qd[j], 0.0, self.limit_max_impulse))
                    self.stats["limit_rows"] += 1
        jj = np.array([r[0] for r in rows])
        sg = np.array([r[1] for r in rows])
        dinv = 1.0 / Minv[jj, jj]
        rhs = np.array([r[2] for r in rows]) * dinv
        lo = np.array([r[3] for r in rows])
        hi = np.array([r[4] for r in rows])
        lam = np.zeros(len(rows))
        dv = np.zeros(n)
        m = len(rows)
        for it in range(self.iterations):
            worst = 0.0
            for k in range(m):
                idx = k if (it & 1) else m - 1 - k
                j = jj[idx]
                d = rhs[idx] - sg[idx] * dv[j] * dinv[idx]
                s = lam[idx] + d
                if s < lo[idx]:
                    d, s = lo[idx] - lam[idx], lo[idx]
                elif s > hi[idx]:
                    d, s = hi[idx] - lam[idx], hi[idx]
                lam[idx] = s
                dv += Minv[:, j] * (sg[idx] * d)
                worst = max(worst, (d / dinv[idx]) ** 2)
            self.stats["sweeps"] += 1
            if worst <= self.threshold:
                break
        self.stats["steps"] += 1
        self.stats["max_residual"] = max(self.stats["max_residual"], math.sqrt(worst))
        self.stats["cap_hits"] += int(np.sum((lam <= lo) & (lo < 0)) + np.sum(lam >= hi))
        b.qd = qd + dv
        b.q = b.q + dt * b.qd
        pris = [abs(b.q[j]) for j in range(n) if b.links[j].joint_type == "prismatic"]
        self.stats["max_prismatic"] = max(self.stats["max_prismatic"], max(pris))
