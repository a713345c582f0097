#!/usr/bin/env python
"""Time the reference's OWN Python env (Environment/Exoskeleton_env.py, unmodified) the way the training script drives it
(Simulation/Exoskeleton_agent_train.py:139-147: env.step per env in a Python loop), BASELINE config 1 settings
(type [1,1,1,1], motion 0, seed 0, evaluation DR settings), U(-1,1)^7 actions, 1000 steps per env, ONE PROCESS PER HOST CORE.

TEST INFRASTRUCTURE ONLY; needs /root/reference, so it runs in the build container, not on the GPU box. pybullet and gym are
not installed (and not installable): the six pybullet calls the env makes are served by oracle/minibullet.py
("real PyBullet: unavailable on this box"); everything else - numpy RNG, scipy solve_ivp RK45 x2 per step, the env's own
Python - is the reference's code. Also times the script-side per-step metric (3 x forward_kinematics, train:158-170).

    python oracle/time_reference_env.py [--procs N] [--steps K] [--motor ideal|full]
"""
import argparse
import multiprocessing as mp
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))


def worker(args):
    rank, steps, motor, with_metric = args
    try:
        os.sched_setaffinity(0, {rank % os.cpu_count()})
    except (AttributeError, OSError):
        pass
    import contextlib
    import io
    import numpy as np
    sys.path.insert(0, HERE)
    import gen_golden as G
    Env = G._import_reference(dict(link_state_lag=1, motor_model=motor))
    from Utilities.calculate_arm_end_effector_points import forward_kinematics, distance_3d
    np.random.seed(rank)
    with contextlib.redirect_stdout(io.StringIO()):
        env = Env(reference_motion_file_num="0", tremor_sequence=np.array([1, 1, 1, 1, 0, 0, 0]), **G.EVAL_SETTINGS)
        env.reset()
        rs = np.random.RandomState(1000 + rank)
        acts = rs.uniform(-1, 1, (steps, 7))
        for t in range(20):                       # warm-up (imports, scipy first call)
            if env.step(acts[t])[2]:
                env.reset()
        t0 = time.perf_counter()
        for t in range(steps):
            obs, r, done, _, info = env.step(acts[t])
            if with_metric:                       # Exoskeleton_agent_train.py:158-170
                ang = np.radians(np.array(env.return_original_joint_angles(), dtype=float))
                p0 = forward_kinematics(ang, 0.4, 0.4, 0.05)
                p1 = forward_kinematics(ang + np.radians(info["ampl_val"]), 0.4, 0.4, 0.05)
                p2 = forward_kinematics(ang + np.radians(info["tremor_ampl_val"]), 0.4, 0.4, 0.05)
                _ = (distance_3d(p1, p0) - distance_3d(p2, p0))
            if done:
                env.reset()
        dt = time.perf_counter() - t0
    return steps / dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--motor", default="ideal")
    args = ap.parse_args()
    for with_metric in (False, True):
        with mp.get_context("spawn").Pool(args.procs) as pool:
            t0 = time.perf_counter()
            rates = pool.map(worker, [(r, args.steps, args.motor, with_metric) for r in range(args.procs)])
            wall = time.perf_counter() - t0
        print("reference ExoskeletonEnv_train on minibullet(%s), %d processes x %d steps%s: %.1f env-steps/s per core "
              "(min %.1f max %.1f), %.1f env-steps/s total; wall %.1f s  [real PyBullet: unavailable on this box]"
              % (args.motor, args.procs, args.steps, " + script-side FK metric" if with_metric else "",
                 sum(rates) / len(rates), min(rates), max(rates), sum(rates), wall))


if __name__ == "__main__":
    main()
