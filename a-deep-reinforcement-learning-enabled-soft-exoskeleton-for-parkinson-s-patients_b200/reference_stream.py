"""Reference-stream mode: episode parameters drawn on the HOST from ONE numpy legacy stream shared by every env and the
warm-up policy, exactly as the reference's training script does (Utilities/seed_setting_.py:9 seeds the global np.random;
Simulation/Exoskeleton_agent_train.py:51-71 builds 8 envs after it, :112 resets them in order, :130 draws the warm-up
actions from the same stream), then uploaded into the resident env (exo_set_field on the prefetched slot) in place of the
per-env device streams. SURVEY.md section 7, hard part 3 / section 8.2 `exo_set_episode_params`.

The batched env's own mode gives every env a PRIVATE MT19937 stream on the device (embarrassingly parallel, bit-exact per
env); this module is the exact-replay mode for a handful of envs, inherently serial because the reference's stream is.
What is restated here is only the ORDER and arithmetic of the reference's draws (Exoskeleton_env.py:193-217,
generate_parkinson_tremor.py:5-73, domain_randomization_anatomical_matrices.py:4-27, Exoskeleton_sim_pybullet.py:98-107);
the step itself still runs in the CUDA kernels. tests/test_gpu_parity.py replays a golden of 8 reference envs built after
one np.random.seed.
"""
import numpy as np

# Utilities/differential_eq_matrices.py:41-69 (seven_by_seven)
I7 = np.array([[0.269, 0, 0, 0.076, 0, 0, -0.014], [0, 0.196, 0.083, 0, -0.002, 0.009, 0], [0, 0.083, 0.079, 0, 0, 0.011, 0],
               [0.076, 0, 0, 0.076, 0, 0, -0.012], [0, -0.002, 0, 0, 0.002, 0, 0], [0, 0.009, 0.011, 0, 0, 0.003, 0],
               [-0.014, 0, 0, -0.012, 0, 0, 0.003]])
D7 = np.array([[0.756, 0.184, 0.020, 0.187, 0, 0, 0], [0.184, 0.383, 0.267, 0, 0, 0, 0], [0.020, 0.267, 0.524, 0, 0, 0, 0],
               [0.187, 0, 0, 0.607, 0, 0, 0], [0, 0, 0, 0, 0.021, 0.001, 0.008], [0, 0, 0, 0, 0.001, 0.028, -0.003],
               [0, 0, 0, 0, 0.008, -0.003, 0.082]])
S7 = np.array([[10.80, 2.626, 0.279, 2.670, 0, 0, 0], [2.626, 5.468, 3.821, 0, 0, 0, 0], [0.279, 3.821, 7.486, 0, 0, 0, 0],
               [2.670, 0, 0, 8.670, 0, 0, 0], [0, 0, 0, 0, 0.756, 0.018, 0.291], [0, 0, 0, 0, 0.018, 0.992, -0.099],
               [0, 0, 0, 0, 0.291, -0.099, 2.920]])
JOINT_MAX = np.array([2.5, 5, 10, 5, 5, 0.5, 0.5])          # generate_parkinson_tremor.py:59


def propagator(I, D, S, dt=1.0 / 40.0):
    """q(dt) = G T for I q'' + D q' + S q = T from rest with constant T (what two scipy RK45 solves per step integrate,
    calculate_joint_angles.py:5-22): G = [1 0] A^-1 (e^{A dt} - 1) [0; I^-1], via the augmented-matrix exponential."""
    from scipy.linalg import expm
    Ii = np.linalg.inv(I)
    M = np.zeros((21, 21))
    M[:7, 7:14] = np.eye(7)
    M[7:14, :7], M[7:14, 7:14], M[7:14, 14:] = -Ii @ S, -Ii @ D, Ii
    return expm(M * dt)[:7, 14:]


class ReferenceStream(object):
    """np.random.seed(seed) followed by the reference's draws, in the reference's order."""

    def __init__(self, seed):
        self.rs = np.random.RandomState(int(seed))

    def constructor(self, length, tremor_sequence, **settings):
        """ExoskeletonEnv_train.__init__: one rand() (Exoskeleton_env.py:70), then initialize_movement (:172)."""
        self.rs.rand()
        return self.initialize_movement(length, tremor_sequence, **settings)

    def initialize_movement(self, length, tremor_sequence, *, tremor_amplitude_range, first_harmonics_interval,
                            second_harmonics_interval, max_force_shoulder, max_force_elbow, dr_actuator_end_pos_shift,
                            dr_actuator_range, matrix_noise_fraction, dt=1.0 / 40.0):
        """Exoskeleton_env.py:193-217: the parameters of one episode of an env whose motion has `length` rows."""
        rs, L = self.rs, int(length)
        mag = tremor_amplitude_range[0] + rs.rand() * (tremor_amplitude_range[1] - tremor_amplitude_range[0])
        f1 = rs.uniform(low=first_harmonics_interval[0], high=first_harmonics_interval[1])
        f2 = rs.uniform(low=second_harmonics_interval[0], high=second_harmonics_interval[1])
        noise = rs.rand(L) * 0.001
        t = np.linspace(0, L * dt, L)
        w1, w2 = np.sin(2 * np.pi * f1 * t), np.sin(2 * np.pi * f2 * t)
        jmax = JOINT_MAX * mag
        table = np.zeros((7, L))
        for j, on in enumerate(tremor_sequence):
            a1 = 10 ** (rs.uniform(low=-5, high=0) / 20)
            a2 = 10 ** (rs.uniform(low=-20, high=-10) / 20)
            g = (a1 * w1 + a2 * w2 + noise) * on
            with np.errstate(invalid="ignore", divide="ignore"):
                g = (-1 + 2 * (g - g.min()) / (g.max() - g.min())) * jmax[j]
            g = np.nan_to_num(g, nan=0, posinf=0, neginf=0)
            table[j] = g * rs.choice([-1, 1], size=L)

        def noisy(M):
            n = rs.uniform(-matrix_noise_fraction, matrix_noise_fraction, size=M.shape)
            return M + (n + n.T) / 2 * M
        I, D, S = noisy(I7), noisy(D7), noisy(S7)
        shift = rs.uniform(low=-dr_actuator_end_pos_shift, high=dr_actuator_end_pos_shift, size=(14, 3))
        fmax_sh = max_force_shoulder * rs.uniform(1 - dr_actuator_range, 1 + dr_actuator_range)
        fmax_el = max_force_elbow * rs.uniform(1 - dr_actuator_range, 1 + dr_actuator_range)
        return dict(tremor_mag=mag, tremor_table=table, I=I, D=D, S=S, shift=shift, force_max=np.array([fmax_sh, fmax_el]),
                    G=propagator(I, D, S, dt))

    def warmup_action(self):
        """Exoskeleton_agent_train.py:130: np.random.uniform(-1, 1, 7) from the same stream."""
        return self.rs.uniform(-1, 1, 7)


def upload_next_episode(env, index, params):
    """Write host-generated parameters into the PREFETCHED slot of env `index` of a BatchedExoskeletonEnv: the next
    env.reset(mask) / auto-reset of that env starts the episode they describe. Read-modify-write of whole fields, so this
    is for a handful of envs (the reference's 8), not for a 65 536-env population."""
    import torch
    L = params["tremor_table"].shape[1]

    def put(name, row):
        cur = env.get_field(name, next=True)
        flat = cur.reshape(env.n, -1)
        flat[index] = torch.as_tensor(np.asarray(row, dtype=np.float64).reshape(-1), device=env.device).to(flat.dtype)
        env.set_field(name, flat, next=True)
    tt = env.get_field("tremor_table", next=True)              # [N, table_len, axes]
    axes = tt.shape[2]
    tt[index] = 0
    tt[index, :L, :] = torch.as_tensor(params["tremor_table"][:axes].T.copy(), device=env.device).to(tt.dtype)
    env.set_field("tremor_table", tt, next=True)
    put("tremor_mag", [params["tremor_mag"]])
    put("IDS", np.concatenate([params["I"].ravel(), params["D"].ravel(), params["S"].ravel()]))
    put("shift", params["shift"])
    put("force_max", params["force_max"])
    put("G", params["G"])
