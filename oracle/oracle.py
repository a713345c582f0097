"""ctypes front-end of the CPU oracle (oracle/exo_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs. The product never imports this module.

Geometry and motion tables come from tests/golden/ref_data_and_units.npz (written by
oracle/gen_golden.py from an independent parse of the reference's exo_v3.urdf and from the
reference's own read_env_texts), so the oracle shares no parsing code with the product.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
LIB = os.path.join(HERE, "libexo_oracle.so")
DATA = os.path.join(REPO, "tests", "golden", "ref_data_and_units.npz")

NL = 19


class OrcGeometry(C.Structure):
    _fields_ = [("parent", C.c_int * NL), ("joint_type", C.c_int * NL),
                ("origin_xyz", (C.c_double * 3) * NL), ("origin_R", (C.c_double * 9) * NL),
                ("axis", (C.c_double * 3) * NL), ("com_xyz", (C.c_double * 3) * NL),
                ("lower", C.c_double * NL), ("upper", C.c_double * NL), ("base_pos", C.c_double * 3)]


class OrcConfig(C.Structure):
    _fields_ = [("tremor_amplitude_range", C.c_double * 2), ("first_harmonics_interval", C.c_double * 2),
                ("second_harmonics_interval", C.c_double * 2),
                ("max_force_shoulder", C.c_double), ("max_force_elbow", C.c_double),
                ("dr_actuator_end_pos_shift", C.c_double), ("dr_actuator_range", C.c_double),
                ("matrix_noise_fraction", C.c_double),
                ("tremor_sequence", C.c_int * 7), ("link_state_lag", C.c_int), ("joint_limits", C.c_int),
                ("ode_mode", C.c_int)]


def build(force=False):
    src = [os.path.join(HERE, "exo_oracle.c"), os.path.join(HERE, "exo_oracle.h")]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in src):
        subprocess.check_call(["make", "-C", HERE, "-s", "-B"])
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        dp, fp, ip = C.POINTER(C.c_double), C.POINTER(C.c_float), C.POINTER(C.c_int)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.POINTER(OrcConfig), C.POINTER(OrcGeometry), dp, C.c_int, C.c_uint32]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_reset.restype = C.c_int
        L.orc_reset.argtypes = [C.c_void_p, fp]
        L.orc_step.argtypes = [C.c_void_p, dp, fp, dp, ip, dp, dp]
        L.orc_get_episode.argtypes = [C.c_void_p, dp, dp, dp, dp, dp, dp, dp]
        L.orc_get_bullet_q.argtypes = [C.c_void_p, dp]
        L.orc_get_counts.restype = C.c_int
        L.orc_get_counts.argtypes = [C.c_void_p]
        L.orc_get_limit_crossings.restype = C.c_long
        L.orc_get_limit_crossings.argtypes = [C.c_void_p]
        L.orc_forward_kinematics.argtypes = [dp, C.c_double, C.c_double, C.c_double, dp]
        L.orc_amplitude_metric.restype = C.c_double
        L.orc_amplitude_metric.argtypes = [C.c_void_p, dp, dp, C.c_double, C.c_double, C.c_double]
        L.orc_solve_diff_eq_rk45.argtypes = [dp, dp, dp, dp, C.c_double, dp, ip]
        L.orc_solve_diff_eq_exact.argtypes = [dp, dp, dp, dp, C.c_double, dp]
        L.orc_propagator.argtypes = [dp, dp, dp, C.c_double, dp]
        L.orc_mt_seed.argtypes = [C.POINTER(C.c_uint32), C.c_uint32]
        L.orc_mt_random_sample.restype = C.c_double
        L.orc_mt_random_sample.argtypes = [C.POINTER(C.c_uint32)]
        L.orc_step_many.argtypes = [C.POINTER(C.c_void_p), C.c_int, dp, fp, dp, ip, C.c_int, C.c_int]
        L.orc_max_threads.restype = C.c_int
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


_data = None


def data():
    global _data
    if _data is None:
        _data = dict(np.load(DATA))
    return _data


def geometry():
    d = data()
    g = OrcGeometry()
    for i in range(NL):
        g.parent[i] = int(d["geo/parent"][i])
        g.joint_type[i] = int(d["geo/joint_type"][i])
        g.lower[i] = float(d["geo/lower"][i])
        g.upper[i] = float(d["geo/upper"][i])
        for k in range(3):
            g.origin_xyz[i][k] = float(d["geo/origin_xyz"][i, k])
            g.axis[i][k] = float(d["geo/axis"][i, k])
            g.com_xyz[i][k] = float(d["geo/com_xyz"][i, k])
        for k in range(9):
            g.origin_R[i][k] = float(d["geo/origin_R"][i, k])
    for k in range(3):
        g.base_pos[k] = float(d["geo/base_pos"][k])
    return g


def motion(m):
    return np.ascontiguousarray(data()[f"motion/{int(m)}"], dtype=np.float64)


MOTION_LENGTHS = [232, 316, 282, 295, 270, 347, 266, 273]

DEFAULTS = dict(tremor_amplitude_range=(0.1, 1.0), first_harmonics_interval=(4.0, 6.0),
                second_harmonics_interval=(8.0, 10.0), max_force_shoulder=40.0, max_force_elbow=20.0,
                dr_actuator_end_pos_shift=0.02, dr_actuator_range=0.1, matrix_noise_fraction=0.1,
                link_state_lag=1, joint_limits=0, ode_mode=0)


def make_config(tremor_sequence, **kw):
    s = dict(DEFAULTS)
    s.update(kw)
    c = OrcConfig()
    for name in ("tremor_amplitude_range", "first_harmonics_interval", "second_harmonics_interval"):
        v = s[name]
        getattr(c, name)[0] = float(v[0])
        getattr(c, name)[1] = float(v[1])
    for name in ("max_force_shoulder", "max_force_elbow", "dr_actuator_end_pos_shift", "dr_actuator_range",
                 "matrix_noise_fraction"):
        setattr(c, name, float(s[name]))
    for j in range(7):
        c.tremor_sequence[j] = int(tremor_sequence[j])
    c.link_state_lag = int(s["link_state_lag"])
    c.joint_limits = int(s["joint_limits"])
    c.ode_mode = int(s["ode_mode"])
    return c


class OracleEnv(object):
    """Same life-cycle as the reference: np.random.seed(seed); ExoskeletonEnv_train(...); reset(); step()."""

    def __init__(self, motion_id, tremor_sequence, seed, **settings):
        self.L = lib()
        self.cfg = make_config(tremor_sequence, **settings)
        self.geo = geometry()
        self.motion_id = int(motion_id)
        mo = motion(motion_id)
        self.len = mo.shape[1]
        self.h = self.L.orc_create(C.byref(self.cfg), C.byref(self.geo), _dp(mo), self.len, int(seed) & 0xFFFFFFFF)
        if not self.h:
            raise RuntimeError("orc_create failed")

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_destroy(self.h)
            self.h = None

    def reset(self):
        obs = np.zeros(80, np.float32)
        score = self.L.orc_reset(self.h, _fp(obs))
        return obs, score

    def step(self, action):
        a = np.ascontiguousarray(action, dtype=np.float64)
        obs = np.zeros(80, np.float32)
        rew = C.c_double()
        done = C.c_int()
        info = np.zeros(35)
        parts = np.zeros(5)
        self.L.orc_step(self.h, _dp(a), _fp(obs), C.byref(rew), C.byref(done), _dp(info), _dp(parts))
        inf = dict(actuator_torques=info[0:7], torque_val=info[7:14], ampl_val=info[14:21],
                   tremor_torque_val=info[21:28], tremor_ampl_val=info[28:35],
                   reward_unwanted=parts[0], reward_torque=parts[1], reward_axis=parts[2],
                   reward_control=parts[3], reward_smoothness=parts[4])
        return obs, rew.value, bool(done.value), False, inf

    def episode(self):
        tt = np.zeros((7, self.len))
        I, D, S = np.zeros((7, 7)), np.zeros((7, 7)), np.zeros((7, 7))
        shift = np.zeros((14, 3))
        fmax = np.zeros(2)
        mag = C.c_double()
        self.L.orc_get_episode(self.h, _dp(tt), _dp(I), _dp(D), _dp(S), _dp(shift), _dp(fmax), C.byref(mag))
        return dict(tt=tt, I=I, D=D, S=S, shift=shift, fmax=fmax, mag=mag.value)

    def bullet_q(self):
        q = np.zeros(5)
        self.L.orc_get_bullet_q(self.h, _dp(q))
        return q

    @property
    def counts(self):
        return self.L.orc_get_counts(self.h)

    @property
    def limit_crossings(self):
        return self.L.orc_get_limit_crossings(self.h)

    def amplitude_metric(self, ampl, tampl, L1=0.4, L2=0.4, L3=0.05):
        a = np.ascontiguousarray(ampl, dtype=np.float64)
        t = np.ascontiguousarray(tampl, dtype=np.float64)
        return self.L.orc_amplitude_metric(self.h, _dp(a), _dp(t), L1, L2, L3)


class OracleBatch(object):
    """n independent oracle envs stepped with host threads (timed CPU baseline)."""

    def __init__(self, motion_ids, sequences, seeds, **settings):
        self.envs = [OracleEnv(m, s, sd, **settings) for m, s, sd in zip(motion_ids, sequences, seeds)]
        self.n = len(self.envs)
        self.handles = (C.c_void_p * self.n)(*[e.h for e in self.envs])
        self.obs = np.zeros((self.n, 80), np.float32)
        self.reward = np.zeros(self.n)
        self.done = np.zeros(self.n, np.int32)

    def reset(self):
        for i, e in enumerate(self.envs):
            self.obs[i] = e.reset()[0]
        return self.obs

    def dephase(self, period, threads=0, seed=7):
        """Spread the envs over their episodes (bench.py does the same on the GPU): env i takes (i mod period) extra steps
        with random actions and auto-reset, so a timed window holds its steady-state share of episode ends."""
        rs = np.random.RandomState(seed)
        phase = np.arange(self.n) % period
        obs, rew, done = np.zeros((self.n, 80), np.float32), np.zeros(self.n), np.zeros(self.n, np.int32)
        for t in range(period):
            idx = np.nonzero(phase > t)[0]
            if idx.size == 0:
                break
            hs = (C.c_void_p * idx.size)(*[self.envs[i].h for i in idx])
            a = np.ascontiguousarray(rs.uniform(-1, 1, (idx.size, 7)))
            lib().orc_step_many(hs, int(idx.size), _dp(a), _fp(obs), _dp(rew), done.ctypes.data_as(C.POINTER(C.c_int)), 1, int(threads))

    def step(self, actions, auto_reset=True, threads=0):
        a = np.ascontiguousarray(actions, dtype=np.float64)
        lib().orc_step_many(self.handles, self.n, _dp(a), _fp(self.obs), _dp(self.reward),
                            self.done.ctypes.data_as(C.POINTER(C.c_int)), int(auto_reset), int(threads))
        return self.obs, self.reward, self.done


def solve_diff_eq_rk45(I, D, S, T, dt=1 / 40):
    q = np.zeros(7)
    n = C.c_int()
    lib().orc_solve_diff_eq_rk45(_dp(np.ascontiguousarray(I, dtype=np.float64)), _dp(np.ascontiguousarray(D, dtype=np.float64)),
                                 _dp(np.ascontiguousarray(S, dtype=np.float64)), _dp(np.ascontiguousarray(T, dtype=np.float64)), dt, _dp(q), C.byref(n))
    return q, n.value


def propagator(I, D, S, dt=1 / 40):
    G = np.zeros((7, 7))
    lib().orc_propagator(_dp(np.ascontiguousarray(I, dtype=np.float64)), _dp(np.ascontiguousarray(D, dtype=np.float64)),
                         _dp(np.ascontiguousarray(S, dtype=np.float64)), dt, _dp(G))
    return G


def forward_kinematics(theta, L1=0.4, L2=0.4, L3=0.05):
    p = np.zeros(3)
    lib().orc_forward_kinematics(_dp(np.ascontiguousarray(theta, dtype=np.float64)), L1, L2, L3, _dp(p))
    return p


def mt_random_samples(seed, n):
    st = (C.c_uint32 * 625)()
    lib().orc_mt_seed(st, int(seed))
    return np.array([lib().orc_mt_random_sample(st) for _ in range(n)])
