#!/usr/bin/env python
"""Prints worst-case differences between the CUDA env and (a) the CPU oracle, (b) the reference goldens.
Diagnostic companion of tests/test_gpu_parity.py (same comparisons, but reports numbers instead of asserting).

    python tools/gpu_parity_report.py [--precision fp32|fp64] [--steps 50]
"""
import argparse
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "oracle"))
import exo_b200  # noqa: E402
import oracle as O  # noqa: E402

TYPES15 = [[(m >> 3) & 1, (m >> 2) & 1, (m >> 1) & 1, m & 1, 0, 0, 0] for m in range(1, 16)]
NODR = dict(tremor_amplitude_range=[1.0, 1.0], first_harmonics_interval=[4.0, 6.0], second_harmonics_interval=[8.0, 10.0],
            max_force_shoulder=40.0, max_force_elbow=20.0, dr_actuator_end_pos_shift=0.0, dr_actuator_range=0.0,
            matrix_noise_fraction=0.0)
TRAIN = dict(tremor_amplitude_range=[0.1, 1.0], first_harmonics_interval=[4.0, 6.0], second_harmonics_interval=[8.0, 10.0],
             max_force_shoulder=40.0, max_force_elbow=20.0, dr_actuator_end_pos_shift=0.02, dr_actuator_range=0.03,
             matrix_noise_fraction=0.1)


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(1e-300, np.max(np.abs(b))))


def run(precision, settings, steps, ode_mode, label, seeds0=100):
    mids, seqs, seeds = [], [], []
    for ti, s in enumerate(TYPES15):
        for m in range(8):
            mids.append(m); seqs.append(s); seeds.append(seeds0 + ti * 8 + m)
    n = len(mids)
    env = exo_b200.BatchedExoskeletonEnv(mids, seqs, seeds, precision=precision, **settings)
    orc = [O.OracleEnv(mids[i], seqs[i], seeds[i], ode_mode=ode_mode, **settings) for i in range(n)]
    obs = env.reset().cpu().numpy()
    obs_o = np.stack([o.reset()[0] for o in orc])
    W = {"obs0": float(np.max(np.abs(obs - obs_o)))}
    # episode parameters
    E = [o.episode() for o in orc]
    ids = env.get_field("IDS").cpu().numpy()
    W["IDS_exact"] = bool(all(np.array_equal(ids[i], np.concatenate([E[i]["I"].ravel(), E[i]["D"].ravel(), E[i]["S"].ravel()])) for i in range(n)))
    W["mag_exact"] = bool(np.array_equal(env.get_field("tremor_mag").cpu().numpy()[:, 0], np.array([e["mag"] for e in E])))
    sh = env.get_field("shift").cpu().numpy()
    fm = env.get_field("force_max").cpu().numpy()
    tt = env.get_field("tremor_table").cpu().numpy()
    real = np.float32 if precision == "fp32" else np.float64
    W["shift_exact"] = bool(all(np.array_equal(sh[i], E[i]["shift"].ravel().astype(real)) for i in range(n)))
    W["fmax_exact"] = bool(all(np.array_equal(fm[i], E[i]["fmax"].astype(real)) for i in range(n)))
    ex, tot, worst = 0, 0, 0.0
    for i in range(n):
        L = orc[i].len
        ref = E[i]["tt"][:tt.shape[2], :L].T.astype(real)
        got = tt[i, :L, :]
        ex += int(np.sum(got == ref)); tot += ref.size
        worst = max(worst, float(np.max(np.abs(got - ref))))
    W["table_exact_frac"] = ex / tot
    W["table_maxabs"] = worst
    G = env.get_field("G").cpu().numpy()
    W["G_rel"] = max(rel(G[i].reshape(7, 7), O.propagator(E[i]["I"], E[i]["D"], E[i]["S"])) for i in range(n))
    rs = np.random.RandomState(7)
    keys = ["actuator_torques", "torque_val", "ampl_val", "tremor_torque_val", "tremor_ampl_val"]
    for k in keys + ["obs", "reward", "qb", "rew_parts", "amp_metric"]:
        W[k] = 0.0
    for t in range(steps):
        a = rs.uniform(-1, 1, (n, 7)).astype(np.float32)
        ob, rw, dn, info = env.step(torch.from_numpy(a).cuda(), with_info=True)
        ob, rw, dn = ob.cpu().numpy(), rw.cpu().numpy(), dn.cpu().numpy()
        inf = {k: v.cpu().numpy() for k, v in info.items()}
        qb = env.get_field("q").cpu().numpy()
        for i in range(n):
            o_ob, o_r, o_d, _, o_inf = orc[i].step(a[i].astype(np.float64))
            assert bool(dn[i]) == o_d
            W["obs"] = max(W["obs"], float(np.max(np.abs(ob[i] - o_ob))))
            W["reward"] = max(W["reward"], abs(float(rw[i]) - o_r))
            W["qb"] = max(W["qb"], float(np.max(np.abs(qb[i] - orc[i].bullet_q()))))
            for k in keys:
                W[k] = max(W[k], rel(inf[k][i], o_inf[k]))
            parts = np.array([o_inf[k] for k in exo_b200.REWARD_KEYS])
            W["rew_parts"] = max(W["rew_parts"], float(np.max(np.abs(inf["reward_parts"][i] - parts))))
            am = orc[i].amplitude_metric(o_inf["ampl_val"], o_inf["tremor_ampl_val"])
            W["amp_metric"] = max(W["amp_metric"], abs(float(inf["amp_metric"][i]) - am) / max(1.0, abs(am)))
    print(label, {k: (v if isinstance(v, bool) else float("%.3g" % v)) for k, v in W.items()})
    env.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=40)
    args = ap.parse_args()
    for prec in ("fp64", "fp32"):
        run(prec, NODR, args.steps, 1, "%s noDR vs oracle(exact ODE)" % prec)
        run(prec, TRAIN, args.steps, 1, "%s trainDR vs oracle(exact ODE)" % prec)
        run(prec, TRAIN, args.steps, 0, "%s trainDR vs oracle(RK45 = reference)" % prec)


if __name__ == "__main__":
    main()
