"""minibullet — CPU restatement of the pybullet boundary of the Exoskeleton env.

TEST INFRASTRUCTURE ONLY (oracle). Nothing in the product imports this file.

What it restates: pybullet 3.2.5 (README.md:58 of the reference; an un-vendored pip
wheel, absent from /root/reference and from this image) as used by
  Environment/Exoskeleton_env.py:113-118,433          (client setup, stepSimulation)
  Environment/Exoskeleton_sim_pybullet.py:18          (loadURDF fixed base at (0,0,0.1))
  Environment/Exoskeleton_sim_pybullet.py:115-117     (setJointMotorControlArray POSITION_CONTROL)
  Environment/Exoskeleton_sim_pybullet.py:129-142,194-195,348-349 (getLinkState(...)[0])

PARITY UNPINNED: the reference has no tests or golden vectors at this boundary and
pybullet cannot be installed here, so the model below is Bullet's published algorithm
restated from its documented behaviour, with knobs for the two points that could not
be verified against a running Bullet:

* Link indices: URDF importer numbers links in pre-order DFS over child joints in file
  order, base = -1 (reproduces every handle hard-coded in
  Exoskeleton_sim_pybullet.py:21-25,48-63).
* getLinkState(...)[0] is the link's *inertial frame* (COM) origin in world coordinates,
  read from the per-link cached world transform. That cache is filled at load time and at
  the forward-kinematics stage at the START of each stepSimulation, i.e. before the joint
  positions are integrated -> reads lag the joint state by one step
  (knob `link_state_lag`: 1 = that behaviour, 0 = cache refreshed after integration).
* POSITION_CONTROL with pybullet's default gains (positionGain kp=0.1, velocityGain kd=1.0,
  targetVelocity 0, maxForce 1e5 N): Bullet's joint motor row asks for the joint velocity
      v* = kp*(target-q)/dt + qdot + kd*(0-qdot)
  subject to |impulse| <= maxForce*dt. For this chain (masses of a few kg, dt=1/40) the cap
  never binds, and at solver convergence qdot+ = v*, then q+ = q + dt*qdot+ (semi-implicit
  Euler). With kd=1 this is q+ = q + kp*(target-q): a first-order tracker, independent of
  gravity, link masses and inertias. Motors persist until re-commanded.
* The 14 prismatic dummy joints carry pybullet's default velocity motors (target 0,
  max impulse 1 N*s per step >> gravity impulse 0.245 N*s of a 1 kg link): they stay at 0.
* Joint limits (exo_v3.urdf:17,37,57,77,97): knob `joint_limits` "off" (default: the motor
  row is taken to win) or "clamp" (q+ clipped into [lower, upper]). The oracle reports how
  often a step would have crossed a limit so the regime can be checked.

`motor_model="full"` (round 2) does not ASSUME that closed form: it restates one Bullet
multibody step for this tree from Bullet's published algorithm (btMultiBodyDynamicsWorld /
btMultiBody / btMultiBodyJointMotor / btMultiBodyJointLimitConstraint /
btMultiBodyConstraintSolver, Bullet 3.2x as wrapped by pybullet 3.2.5), so that the closed form
is a RESULT whose error can be stated (tests/test_multibody_model.py):

  1. forward kinematics of the 19-DoF tree (5 revolute + 14 prismatic); link-state cache refresh
  2. external forces: gravity m*g on every link COM, Bullet's per-link velocity damping
     (linear/angular 0.04: force -m*v*(k+k|v|), torque -I*w*(k+k|w|)), gyroscopic term w x Iw
  3. unconstrained joint accelerations (articulated-body algorithm == M(q)^-1 (tau - bias); here
     via composite mass matrix + recursive Newton-Euler bias, same result to rounding),
     qd += dt*qdd
  4. constraint rows, in the order Bullet's (unstable) quickSort of the world's constraint list
     [limit_0..limit_18, motor_0..motor_18] leaves them (all in one island):
       * motor row per joint: Jacobian e_j, desired velocity kp*(target-q)/dt + qd + kd*(vt-qd)
         (position motors on joints 0-4: kp 0.1, kd 1, impulse cap maxForce*dt = 2500;
          pybullet's default velocity motors on the 14 prismatic joints: kp 0, kd 1, target 0,
          impulse cap 1.0), rhs = (desired - J.qd) / (J M^-1 J^T)
       * joint-limit rows only while q is at/over a bound: unilateral, impulse in [0, 100],
         rhs = (erp*(-penetration)/dt - J.qd) / (J M^-1 J^T), erp 0.2
  5. projected Gauss-Seidel ("sequential impulse"): numIterations 50 (pybullet default), sweep
     direction reversed on even iterations, early exit when the largest squared row velocity
     error of a sweep is <= leastSquaresResidualThreshold (pybullet sets 1e-7)
  6. q += dt*qd (semi-implicit Euler); no contacts (no self-collision flag, no ground)
  Link inertias: pybullet recomputes them from the collision boxes unless
  URDF_USE_INERTIA_FROM_FILE is passed (the reference passes no flags,
  Exoskeleton_sim_pybullet.py:18) -> knob `inertia_from` "collision" (default) / "file".
All of it is [bullet-memory]: restated from the published algorithm, not checked against a running
Bullet (pybullet is not installable here or on the GPU box, profiles/r02a_pybullet_probe.txt).
"""
import math
import xml.etree.ElementTree as ET

import numpy as np


def rpy_matrix(r, p, y):
    """URDF fixed-axis roll/pitch/yaw -> rotation matrix Rz(y) @ Ry(p) @ Rx(r)."""
    cr, sr = math.cos(r), math.sin(r)
    cp, sp = math.cos(p), math.sin(p)
    cy, sy = math.cos(y), math.sin(y)
    rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]], dtype=np.float64)
    ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]], dtype=np.float64)
    rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]], dtype=np.float64)
    return rz @ ry @ rx


def _vec(s, default=(0.0, 0.0, 0.0)):
    if s is None:
        return np.array(default, dtype=np.float64)
    return np.array([float(x) for x in s.split()], dtype=np.float64)


class Link(object):
    __slots__ = ("name", "com_xyz", "parent", "joint_type", "axis", "origin_xyz", "origin_R",
                 "lower", "upper", "children", "mass", "inertia_file", "R_inertial", "boxes")


class Body(object):
    """Kinematic tree of one URDF with Bullet's link numbering (base = -1)."""

    @classmethod
    def from_arrays(cls, d):
        """Rebuild the tree from the `geo/*` arrays gen_golden.py stored in tests/golden/ref_data_and_units.npz
        (written by __init__ below from the reference's exo_v3.urdf), so tests run without /root/reference."""
        self = cls.__new__(cls)
        self.base_name = "robot_base"
        self.base_com = np.zeros(3)
        self.base_pos = np.array(d["geo/base_pos"], dtype=np.float64)
        self.links, self.index_of = [], {self.base_name: -1}
        for i in range(len(d["geo/parent"])):
            ln = Link()
            ln.name = str(d["geo/names"][i])
            ln.com_xyz = np.array(d["geo/com_xyz"][i], dtype=np.float64)
            ln.parent = int(d["geo/parent"][i])
            ln.joint_type = "revolute" if int(d["geo/joint_type"][i]) == 0 else "prismatic"
            ln.axis = np.array(d["geo/axis"][i], dtype=np.float64)
            ln.origin_xyz = np.array(d["geo/origin_xyz"][i], dtype=np.float64)
            ln.origin_R = np.array(d["geo/origin_R"][i], dtype=np.float64).reshape(3, 3)
            ln.lower, ln.upper = float(d["geo/lower"][i]), float(d["geo/upper"][i])
            ln.mass = float(d["geo/mass"][i])
            ln.inertia_file = np.array(d["geo/inertia_file"][i], dtype=np.float64)
            ln.R_inertial = np.array(d["geo/R_inertial"][i], dtype=np.float64).reshape(3, 3)
            ln.boxes = [(np.array(d["geo/box_size"][i], dtype=np.float64), np.array(d["geo/box_xyz"][i], dtype=np.float64),
                         np.array(d["geo/box_R"][i], dtype=np.float64).reshape(3, 3))]
            self.index_of[ln.name] = i
            self.links.append(ln)
        self.n = len(self.links)
        self.q, self.qd, self.targets = np.zeros(self.n), np.zeros(self.n), {}
        self.cache = self.forward_kinematics(self.q)
        return self

    def __init__(self, path, base_position):
        root = ET.parse(path).getroot()
        links = {}
        self.inertial = {}       # name -> (mass, diag inertia[3], inertial-frame rotation, collision boxes)
        for le in root.findall("link"):
            com = (0.0, 0.0, 0.0)
            ine = le.find("inertial")
            if ine is not None and ine.find("origin") is not None:
                com = _vec(ine.find("origin").get("xyz"))
            links[le.get("name")] = np.array(com, dtype=np.float64)
            mass, diag, Rin = 0.0, np.zeros(3), np.eye(3)
            if ine is not None:
                mass = float(ine.find("mass").get("value"))
                it = ine.find("inertia")
                assert all(float(it.get(k, 0.0)) == 0.0 for k in ("ixy", "ixz", "iyz")), "principal axes expected"
                diag = np.array([float(it.get(k)) for k in ("ixx", "iyy", "izz")])
                if ine.find("origin") is not None:
                    Rin = rpy_matrix(*_vec(ine.find("origin").get("rpy")))
            boxes = []
            for ce in le.findall("collision"):
                org = ce.find("origin")
                boxes.append((_vec(ce.find("geometry").find("box").get("size")),
                              _vec(org.get("xyz") if org is not None else None),
                              rpy_matrix(*_vec(org.get("rpy") if org is not None else None))))
            self.inertial[le.get("name")] = (mass, diag, Rin, boxes)
        joints = []
        child_names = set()
        for je in root.findall("joint"):
            org = je.find("origin")
            lim = je.find("limit")
            joints.append(dict(
                type=je.get("type"),
                parent=je.find("parent").get("link"),
                child=je.find("child").get("link"),
                xyz=_vec(org.get("xyz") if org is not None else None),
                rpy=_vec(org.get("rpy") if org is not None else None),
                axis=_vec(je.find("axis").get("xyz") if je.find("axis") is not None else None, (1, 0, 0)),
                lower=float(lim.get("lower")) if lim is not None else -math.inf,
                upper=float(lim.get("upper")) if lim is not None else math.inf,
            ))
            child_names.add(je.find("child").get("link"))
        base = [n for n in links if n not in child_names]
        assert len(base) == 1, base
        self.base_name = base[0]
        self.base_pos = np.array(base_position, dtype=np.float64)
        self.base_com = links[self.base_name]
        self.links = []          # Bullet link index -> Link
        self.index_of = {self.base_name: -1}

        def visit(parent_name):
            for j in joints:  # file order
                if j["parent"] != parent_name:
                    continue
                ln = Link()
                ln.name = j["child"]
                ln.com_xyz = links[ln.name]
                ln.parent = self.index_of[parent_name]
                ln.joint_type = j["type"]
                ln.axis = j["axis"]
                ln.origin_xyz = j["xyz"]
                ln.origin_R = rpy_matrix(*j["rpy"])
                ln.lower, ln.upper = j["lower"], j["upper"]
                ln.mass, ln.inertia_file, ln.R_inertial, ln.boxes = self.inertial[ln.name]
                self.index_of[ln.name] = len(self.links)
                self.links.append(ln)
                visit(ln.name)

        visit(self.base_name)
        self.n = len(self.links)
        self.q = np.zeros(self.n)
        self.qd = np.zeros(self.n)
        self.targets = {}        # joint index -> target position (POSITION_CONTROL motors)
        self.cache = self.forward_kinematics(self.q)   # loadURDF fills the link transform cache

    def forward_kinematics(self, q):
        """World COM position of every link for joint vector q (list of 3-vectors)."""
        # loadURDF's basePosition places the base inertial frame; link frames hang off the base *link* frame
        R = [None] * self.n
        p = [None] * self.n
        com = [None] * self.n
        base_R = np.eye(3)
        base_p = self.base_pos - base_R @ self.base_com
        for i, ln in enumerate(self.links):
            pR, pp = (base_R, base_p) if ln.parent < 0 else (R[ln.parent], p[ln.parent])
            Ri = pR @ ln.origin_R
            pi = pp + pR @ ln.origin_xyz
            if ln.joint_type == "revolute" or ln.joint_type == "continuous":
                a = ln.axis / np.linalg.norm(ln.axis)
                c, s = math.cos(q[i]), math.sin(q[i])
                K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
                Ri = Ri @ (np.eye(3) + s * K + (1 - c) * (K @ K))
            elif ln.joint_type == "prismatic":
                pi = pi + Ri @ (ln.axis * q[i])
            R[i], p[i] = Ri, pi
            com[i] = pi + Ri @ ln.com_xyz
        return com


class World(object):
    def __init__(self, link_state_lag=1, joint_limits="off", kp=0.1, kd=1.0, motor_model="ideal", **full_knobs):
        assert link_state_lag in (0, 1)
        assert joint_limits in ("off", "clamp")
        assert motor_model in ("ideal", "full")
        self.link_state_lag = link_state_lag
        self.joint_limits = joint_limits
        self.kp, self.kd = kp, kd
        self.motor_model = motor_model
        self.full_knobs = full_knobs     # inertia_from, iterations, residual_threshold, row_order, ... (multibody.FullStep)
        self.full = []
        self.dt = 1.0 / 240.0
        self.gravity = (0.0, 0.0, 0.0)
        self.bodies = []
        self.limit_crossings = 0

    def reset_simulation(self):
        self.bodies = []
        self.full = []

    def load_urdf(self, path, base_position):
        self.bodies.append(Body(path, base_position))
        self.full.append(None)
        return len(self.bodies) - 1

    def set_position_targets(self, body, joints, targets):
        b = self.bodies[body]
        for j, t in zip(joints, targets):
            b.targets[j] = t

    def link_com_world(self, body, link):
        return self.bodies[body].cache[link]

    def step_simulation(self):
        for bi, b in enumerate(self.bodies):
            if self.link_state_lag == 1:
                b.cache = b.forward_kinematics(b.q)        # FK stage precedes integration
            if self.motor_model == "full":
                if self.full[bi] is None:                  # dt / gravity are set after loadURDF by the env's constructor
                    import multibody
                    self.full[bi] = multibody.FullStep(b, self.dt, self.gravity, self.kp, self.kd, **self.full_knobs)
                fs = self.full[bi]
                fs.dt, fs.gravity = self.dt, self.gravity
                for j, t in b.targets.items():             # same crossing statistic as the ideal model
                    qn = b.q[j] + self.kp * (t - b.q[j])
                    if qn < b.links[j].lower or qn > b.links[j].upper:
                        self.limit_crossings += 1
                fs.step()
                if self.link_state_lag == 0:
                    b.cache = b.forward_kinematics(b.q)
                continue
            for j, t in b.targets.items():
                v = self.kp * (t - b.q[j]) / self.dt + b.qd[j] + self.kd * (0.0 - b.qd[j])
                qn = b.q[j] + self.dt * v
                ln = b.links[j]
                if qn < ln.lower or qn > ln.upper:
                    self.limit_crossings += 1
                    if self.joint_limits == "clamp":
                        qn = min(max(qn, ln.lower), ln.upper)
                        v = (qn - b.q[j]) / self.dt
                b.qd[j] = v
                b.q[j] = qn
            if self.link_state_lag == 0:
                b.cache = b.forward_kinematics(b.q)
