#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the reference's OWN Python, unmodified.

TEST INFRASTRUCTURE ONLY. Run in the build container (needs /root/reference):

    python oracle/gen_golden.py            # writes tests/golden/ref_*.npz

What runs: `Environment.Exoskeleton_env.ExoskeletonEnv_train` and every `Utilities/*` module
it imports, straight from /root/reference, with numpy's global legacy RNG and scipy's
solve_ivp exactly as the reference uses them. Only the third-party imports that are absent
here are shimmed (oracle/stubs; pybullet -> oracle/minibullet.py, see its header:
PARITY UNPINNED at the pybullet boundary).

Seeding convention of every case (it is the reference's own: seed_setting_.py:9 then the
constructor): `np.random.seed(seed)`; construct the env (constructor draws one rand(),
Exoskeleton_env.py:70, and runs initialize_movement once, :172); then `reset()` and step.
Actions come from a separate RandomState so the env's stream is not consumed by the policy
(except case family A, which uses the shipped TD7 checkpoint actor and needs no draws).

Families
  A  BASELINE config 1: motion 0, type [1,1,1,1], seed 0, evaluation settings
     (Evaluate_control_performance.py:30-37,46-47), 1000 steps across episode boundaries,
     actions from the shipped [1,1,1,1] checkpoint actor+encoder
     (TD7_multi_agent_Pink_noise.py:208-228, use_checkpoint=True, use_exploration=False).
  B  BASELINE config 2: 15 tremor types x 8 motions, no domain randomisation, 30 steps.
  C  BASELINE config 3 settings (training DR, Exoskeleton_agent_train.py:30-35,43-44),
     one env per motion: a full episode, reset, 20 more steps (state carried across reset).
  S  the training script's SHARED stream (seed_setting_.py:9, Exoskeleton_agent_train.py:51-71,110-147): np.random.seed once,
     8 envs built one after the other (one per motion), reset in order, then lock-step steps whose warm-up actions
     np.random.uniform(-1, 1, 7) come from the same global stream, across the first episode end of the shortest motion.
     Written to ref_shared_stream.npz; replayed through host-generated parameters (exo_b200.reference_stream).
  F  the same env on minibullet's motor_model="full" (the general Bullet multibody step, oracle/multibody.py, pybullet's
     default solver settings): config 1 for 300 steps (one reset) and config 3 on motions 1 and 5. Written to
     ref_full_motor.npz; what the product's EXO_MOTOR_FULL path is held to.
"""
import argparse
import contextlib
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF = os.environ.get("EXO_REFERENCE", "/root/reference")


def _import_reference(knobs):
    sys.path.insert(0, os.path.join(HERE, "stubs"))
    sys.path.insert(0, REF)
    os.chdir(os.path.join(REF, "Simulation"))   # the reference opens files relative to cwd
    import pybullet_utils.bullet_client as bc
    bc.DEFAULT_KNOBS.update(knobs)
    from Environment.Exoskeleton_env import ExoskeletonEnv_train
    return ExoskeletonEnv_train


INFO_KEYS = ["actuator_torques", "torque_val", "ampl_val", "tremor_torque_val", "tremor_ampl_val"]
REW_KEYS = ["reward_unwanted", "reward_torque", "reward_axis", "reward_control", "reward_smoothness"]


class Recorder(object):
    """Accumulates one case: episode parameter snapshots + per-step records."""

    def __init__(self, env):
        self.env = env
        self.ep = dict(tt=[], I=[], D=[], S=[], shift=[], fmax=[], mag=[], obs0=[], at_step=[])
        self.st = dict(action=[], obs=[], reward=[], done=[], qb=[], rew_parts=[], q19=[], qd19=[])
        for k in INFO_KEYS:
            self.st[k] = []

    def snap_episode(self, obs0):
        e = self.env
        self.ep["tt"].append(np.array(e.tremor_torque_values, dtype=np.float64))
        self.ep["I"].append(np.array(e.I_current))
        self.ep["D"].append(np.array(e.D_current))
        self.ep["S"].append(np.array(e.S_current))
        self.ep["shift"].append(np.array(e.exoskeleton_sim_model.dummy_shift_coordinates))
        self.ep["fmax"].append(np.array([e.max_output_shoulder, e.max_output_elbow], dtype=np.float64))
        self.ep["mag"].append(float(e.tremor_magnitude))
        self.ep["obs0"].append(np.array(obs0, dtype=np.float32))
        self.ep["at_step"].append(len(self.st["reward"]))

    def snap_step(self, action, out):
        obs, reward, done, trunc, info = out
        assert trunc is False
        self.st["action"].append(np.array(action, dtype=np.float64))
        self.st["obs"].append(np.array(obs, dtype=np.float32))
        self.st["reward"].append(float(reward))
        self.st["done"].append(bool(done))
        self.st["qb"].append(np.array(self.env.client._world.bodies[0].q[:5]))
        self.st["q19"].append(np.array(self.env.client._world.bodies[0].q))
        self.st["qd19"].append(np.array(self.env.client._world.bodies[0].qd))
        self.st["rew_parts"].append(np.array([float(info[k]) for k in REW_KEYS]))
        for k in INFO_KEYS:
            self.st[k].append(np.array(info[k], dtype=np.float64))

    def pack(self, prefix, out, meta, keep_q19=False):
        for k, v in self.ep.items():
            out[f"{prefix}/ep_{k}"] = np.array(v)
        for k, v in self.st.items():
            if k in ("q19", "qd19") and not keep_q19:
                continue
            out[f"{prefix}/{k}"] = np.array(v)
        for k, v in meta.items():
            out[f"{prefix}/meta_{k}"] = np.array(v)


def f32_actions(rs):
    """U(-1,1)^7 rounded to float32 and widened back (the boundary's action dtype is float32)."""
    return rs.uniform(-1.0, 1.0, 7).astype(np.float32).astype(np.float64)


def run_case(Env, seed, motion, seq, settings, n_steps, policy, quiet=True):
    """settings: dict of ctor kwargs. policy(obs, t) -> action[7] float64. Steps across resets."""
    np.random.seed(seed)
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink if quiet else sys.stdout):
        env = Env(reference_motion_file_num=str(motion), tremor_sequence=np.array(seq), **settings)
        rec = Recorder(env)
        obs, score = env.reset()
        assert score == 2
        rec.snap_episode(obs)
        for t in range(n_steps):
            a = policy(obs, t)
            out = env.step(a)
            rec.snap_step(a, out)
            obs = out[0]
            if out[2]:
                obs, _ = env.reset()
                rec.snap_episode(obs)
    meta = dict(seed=seed, motion=motion, seq=np.array(seq, dtype=np.int64),
                max_count=env.max_count, max_reward=float(env.max_reward),
                limit_crossings=env.client._world.limit_crossings,
                **{k: np.array(v, dtype=np.float64) for k, v in settings.items()})
    return rec, meta


EVAL_SETTINGS = dict(tremor_amplitude_range=[0.95, 1.05], first_harmonics_interval=[3.75, 6.25],
                     second_harmonics_interval=[7.5, 12.5], max_force_shoulder=40.0, max_force_elbow=20.0,
                     dr_actuator_end_pos_shift=0.025, dr_actuator_range=0.04, matrix_noise_fraction=0.125)
NODR_SETTINGS = dict(tremor_amplitude_range=[1.0, 1.0], first_harmonics_interval=[4.0, 6.0],
                     second_harmonics_interval=[8.0, 10.0], max_force_shoulder=40.0, max_force_elbow=20.0,
                     dr_actuator_end_pos_shift=0.0, dr_actuator_range=0.0, matrix_noise_fraction=0.0)
TRAIN_SETTINGS = dict(tremor_amplitude_range=[0.1, 1.0], first_harmonics_interval=[4.0, 6.0],
                      second_harmonics_interval=[8.0, 10.0], max_force_shoulder=40.0, max_force_elbow=20.0,
                      dr_actuator_end_pos_shift=0.02, dr_actuator_range=0.03, matrix_noise_fraction=0.1)

TYPES15 = [[(m >> 3) & 1, (m >> 2) & 1, (m >> 1) & 1, m & 1, 0, 0, 0] for m in range(1, 16)]


def trained_actor_policy():
    """Shipped checkpoint actor for type [1,1,1,1], evaluated the way the reference's
    select_action(use_checkpoint=True, use_exploration=False) does."""
    import torch
    from Agent.TD7_multi_agent_Pink_noise import Actor, Encoder
    import torch.nn.functional as F
    d = os.path.join(REF, "Simulation", "AGENT_NNS", "[1,1,1,1]", "[1,1,1,1]_")
    actor = Actor(80, 7, 300, 300, F.relu)
    enc = Encoder(80, 7, 300, 300, F.elu)
    actor.load_state_dict(torch.load(d + "checkpoint_actor", map_location="cpu", weights_only=False))
    enc.load_state_dict(torch.load(d + "checkpoint_encoder", map_location="cpu", weights_only=False))
    torch.set_num_threads(1)

    def policy(obs, t):
        with torch.no_grad():
            s = torch.from_numpy(np.asarray(obs)[None]).float()
            a = actor(s, enc.zs(s)).cpu().data.numpy()[0]
        return a.clip(-1, 1).astype(np.float64)      # float32 values widened
    return policy


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(REPO, "tests", "golden"))
    ap.add_argument("--lag", type=int, default=1)
    ap.add_argument("--families", default="DABCSF")
    args = ap.parse_args()
    out_dir = os.path.abspath(args.out)
    Env = _import_reference(dict(link_state_lag=args.lag))
    os.makedirs(out_dir, exist_ok=True)

    if "D" in args.families:
        # data tables + unit known-answers, all produced by the reference's own functions
        import minibullet
        from Utilities.read_txt_env import read_env_texts
        from Utilities.calculate_joint_angles import solve_diff_eq
        from Utilities.differential_eq_matrices import seven_by_seven
        from Utilities.calculate_arm_end_effector_points import forward_kinematics
        from scipy.integrate import solve_ivp
        out = {}
        body = minibullet.Body(os.path.join(REF, "Simulation", "exo_v3.urdf"), [0, 0, 0.1])
        out["geo/parent"] = np.array([l.parent for l in body.links], dtype=np.int32)
        out["geo/joint_type"] = np.array([0 if l.joint_type == "revolute" else 1 for l in body.links], dtype=np.int32)
        out["geo/origin_xyz"] = np.array([l.origin_xyz for l in body.links])
        out["geo/origin_R"] = np.array([l.origin_R.reshape(9) for l in body.links])
        out["geo/axis"] = np.array([l.axis for l in body.links])
        out["geo/com_xyz"] = np.array([l.com_xyz for l in body.links])
        out["geo/lower"] = np.array([l.lower for l in body.links])
        out["geo/upper"] = np.array([l.upper for l in body.links])
        out["geo/base_pos"] = np.array(body.base_pos - body.base_com)
        out["geo/names"] = np.array([l.name for l in body.links])
        out["geo/zero_pose_com"] = np.array(body.cache)
        # dynamics data of the same URDF (masses, principal inertias, inertial-frame rotations, collision boxes):
        # what minibullet's motor_model="full" (oracle/multibody.py) needs; exo_v3.urdf:3-13,22-33,...
        out["geo/mass"] = np.array([l.mass for l in body.links])
        out["geo/inertia_file"] = np.array([l.inertia_file for l in body.links])
        out["geo/R_inertial"] = np.array([l.R_inertial.reshape(9) for l in body.links])
        assert all(len(l.boxes) == 1 for l in body.links)
        out["geo/box_size"] = np.array([l.boxes[0][0] for l in body.links])
        out["geo/box_xyz"] = np.array([l.boxes[0][1] for l in body.links])
        out["geo/box_R"] = np.array([l.boxes[0][2].reshape(9) for l in body.links])
        keys = ["elbow_joint_y_positions", "elbow_joint_z_positions", "shoulder_joint_x_positions",
                "shoulder_joint_y_positions", "shoulder_joint_z_positions"]
        for m in range(8):
            d = read_env_texts(f"reference_motions/ref_motion_{m}.txt")
            out[f"motion/{m}"] = np.array([d[k] for k in keys])
        I, D, S = seven_by_seven()
        out["mat/I"], out["mat/D"], out["mat/S"] = I, D, S
        rs = np.random.RandomState(7)
        Ts, qs, nf = [], [], []
        for i in range(24):
            T = rs.uniform(-10, 10, 7) * (rs.rand(7) < 0.7)
            nfev = [0]
            Ts.append(T)
            qs.append(solve_diff_eq(np.zeros(14), [0, 1 / 40], I, D, S, T))
            def fun(t, y, T=T):
                return np.concatenate((y[7:], np.linalg.solve(I, T - D @ y[7:] - S @ y[:7])))
            nf.append(solve_ivp(fun, [0, 1 / 40], np.zeros(14)).nfev)
        out["ode/T"], out["ode/q"], out["ode/nfev"] = np.array(Ts), np.array(qs), np.array(nf)
        th = [np.radians(v) for v in ([0] * 7, [90, 0, 0, 0, 0, 0, 0], [0, 90, 0, 0, 0, 0, 0], [0, 0, 90, 0, 0, 0, 0],
                                      [0, 0, 0, 90, 0, 0, 0], [4, 5, 6, 0, 0, 0, 0], [10, -20, 30, 40, -50, 5, -5])]
        out["fk/theta"] = np.array(th)
        out["fk/p"] = np.array([forward_kinematics(t, 0.4, 0.4, 0.05) for t in th])
        np.savez_compressed(os.path.join(out_dir, "ref_data_and_units.npz"), **out)
        print("D: geometry, motions, ode, fk")

    if "A" in args.families:
        out = {}
        rec, meta = run_case(Env, 0, 0, [1, 1, 1, 1, 0, 0, 0], EVAL_SETTINGS, 1000, trained_actor_policy())
        rec.pack("A0", out, meta)
        np.savez_compressed(os.path.join(out_dir, "ref_cfg1_trained_actor.npz"), **out)
        print("A: return/step", np.mean(rec.st["reward"]), "limit crossings", meta["limit_crossings"])

    if "B" in args.families:
        out = {}
        n = 0
        for ti, seq in enumerate(TYPES15):
            for m in range(8):
                case = ti * 8 + m
                rs = np.random.RandomState(50_000 + case)
                rec, meta = run_case(Env, 100 + case, m, seq, NODR_SETTINGS, 30, lambda o, t: f32_actions(rs))
                rec.pack(f"B{case}", out, meta)
                n += meta["limit_crossings"]
        np.savez_compressed(os.path.join(out_dir, "ref_cfg2_types_x_motions.npz"), **out)
        print("B: 120 cases, limit crossings", n)

    if "C" in args.families:
        out = {}
        n = 0
        for m in range(8):
            seq = TYPES15[(5 * m + 14) % 15]
            rs = np.random.RandomState(60_000 + m)
            L = [232, 316, 282, 295, 270, 347, 266, 273][m]
            rec, meta = run_case(Env, 900 + m, m, seq, TRAIN_SETTINGS, (L - 3) + 20, lambda o, t: f32_actions(rs))
            assert meta["max_count"] == L
            rec.pack(f"C{m}", out, meta)
            n += meta["limit_crossings"]
        np.savez_compressed(os.path.join(out_dir, "ref_cfg3_dr_episodes.npz"), **out)
        print("C: 8 cases, limit crossings", n)


    if "S" in args.families:
        out = {}
        seed, n_steps = 5, 260
        seqs = [TYPES15[(3 * m + 14) % 15] for m in range(8)]
        np.random.seed(seed)                                   # set_seeds (seed_setting_.py:9)
        sink = io.StringIO()
        with contextlib.redirect_stdout(sink):
            envs = [Env(reference_motion_file_num=str(m), tremor_sequence=np.array(seqs[m]), **TRAIN_SETTINGS) for m in range(8)]
            recs = [Recorder(e) for e in envs]
            done = [False] * 8
            obs = [None] * 8
            for i, e in enumerate(envs):                       # train:111-113
                obs[i], _ = e.reset()
                recs[i].snap_episode(obs[i])
            for t in range(n_steps):
                if all(done):                                  # train:123 -> next outer iteration resets every env
                    for i, e in enumerate(envs):
                        obs[i], _ = e.reset()
                        recs[i].snap_episode(obs[i])
                    done = [False] * 8
                acts = [np.random.uniform(-1, 1, 7) if not done[i] else None for i in range(8)]   # train:125-131: one draw per live env
                for i, e in enumerate(envs):
                    if not done[i]:                            # train:139
                        a32 = acts[i].astype(np.float32).astype(np.float64)
                        o = e.step(a32)
                        recs[i].snap_step(a32, o)
                        obs[i], done[i] = o[0], o[2]
        for i in range(8):
            meta = dict(seed=seed, motion=i, seq=np.array(seqs[i], dtype=np.int64), max_count=envs[i].max_count,
                        max_reward=float(envs[i].max_reward), limit_crossings=envs[i].client._world.limit_crossings,
                        **{k: np.array(v, dtype=np.float64) for k, v in TRAIN_SETTINGS.items()})
            recs[i].pack(f"S{i}", out, meta)
        out["S/n_steps"] = np.array(n_steps)
        np.savez_compressed(os.path.join(out_dir, "ref_shared_stream.npz"), **out)
        print("S: 8 envs on one shared stream,", n_steps, "lock-step rounds; steps per env", [len(r.st["reward"]) for r in recs])

    if "F" in args.families:
        import pybullet_utils.bullet_client as bc
        bc.DEFAULT_KNOBS.update(dict(motor_model="full"))
        out = {}
        rec, meta = run_case(Env, 0, 0, [1, 1, 1, 1, 0, 0, 0], EVAL_SETTINGS, 300, trained_actor_policy())
        rec.pack("F0", out, meta, keep_q19=True)
        sweeps = [rec.env.client._world.full[0].stats["sweeps"] / 300.0]
        for k, m in enumerate((1, 5)):
            seq = TYPES15[(5 * m + 14) % 15]
            rs = np.random.RandomState(60_000 + m)
            L = [232, 316, 282, 295, 270, 347, 266, 273][m]
            rec, meta = run_case(Env, 900 + m, m, seq, TRAIN_SETTINGS, (L - 3) + 20, lambda o, t: f32_actions(rs))
            rec.pack(f"F{k + 1}", out, meta, keep_q19=True)
            sweeps.append(rec.env.client._world.full[0].stats["sweeps"] / float((L - 3) + 20))
        bc.DEFAULT_KNOBS.pop("motor_model")
        np.savez_compressed(os.path.join(out_dir, "ref_full_motor.npz"), **out)
        print("F: 3 cases on the general multibody step, sweeps/step", sweeps)


if __name__ == "__main__":
    main()
