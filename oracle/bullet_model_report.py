#!/usr/bin/env python
"""How far is the closed-form position tracker (q+ = q + kp (target - q), product default, minibullet "ideal") from
the general Bullet multibody step (oracle/multibody.py, minibullet "full") on the golden families?

TEST INFRASTRUCTURE ONLY; needs /root/reference (build container). Runs the reference's own Python env
(Environment/Exoskeleton_env.py, unmodified) closed loop on both motor models with identical seeds and actions,
for BASELINE config 1 (family A), config 2 (family B) and config 3 (family C) of oracle/gen_golden.py, and for
every unverifiable choice of the restatement (knobs). Writes the table DESIGN.md section 4 quotes.

    python oracle/bullet_model_report.py [--quick] > profiles/r02_bullet_model_report.txt
"""
import argparse
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import gen_golden as G  # noqa: E402


def run(Env, bc, knobs, *case):
    bc.DEFAULT_KNOBS.clear()
    bc.DEFAULT_KNOBS.update(dict(link_state_lag=1, joint_limits="off"))
    bc.DEFAULT_KNOBS.update(knobs)
    rec, meta = G.run_case(Env, *case)
    return rec, rec.env.client._world


def cases(quick):
    out = []
    # A: config 1, shipped actor, 1000 steps over 4+ episodes
    out.append(("A cfg1 trained actor", [(0, 0, [1, 1, 1, 1, 0, 0, 0], G.EVAL_SETTINGS, 300 if quick else 1000, "actor")]))
    b = []
    for ti, seq in enumerate(G.TYPES15):
        for m in range(8):
            case = ti * 8 + m
            if quick and case % 12:
                continue
            b.append((100 + case, m, seq, G.NODR_SETTINGS, 30, 50_000 + case))
    out.append(("B cfg2 15 types x 8 motions", b))
    c = []
    for m in range(8):
        if quick and m % 4:
            continue
        L = [232, 316, 282, 295, 270, 347, 266, 273][m]
        c.append((900 + m, m, G.TYPES15[(5 * m + 14) % 15], G.TRAIN_SETTINGS, (L - 3) + 20, 60_000 + m))
    out.append(("C cfg3 training DR, full episode + reset", c))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    Env = G._import_reference(dict(link_state_lag=1))
    import pybullet_utils.bullet_client as bc
    actor = G.trained_actor_policy()
    variants = [
        ("full: pybullet defaults (inertia from collision boxes, 50 iterations, residual threshold 1e-7, Bullet row order)", dict(motor_model="full")),
        ("full: URDF_USE_INERTIA_FROM_FILE inertias", dict(motor_model="full", inertia_from="file")),
        ("full: rows in world order (stable sort)", dict(motor_model="full", row_order="natural")),
        ("full: no early exit (threshold 0, always 50 sweeps)", dict(motor_model="full", residual_threshold=0.0)),
        ("full: 1000 iterations, threshold 0 (solver converged)", dict(motor_model="full", residual_threshold=0.0, iterations=1000)),
        ("full: no link damping", dict(motor_model="full", linear_damping=0.0, angular_damping=0.0)),
    ]
    if args.quick:
        variants = variants[:2] + variants[4:5]
    print("closed-form tracker (ideal) vs general multibody step (full), reference env closed loop, same seeds/actions")
    print("columns: max |dq| joint angle (rad) | max |d obs| | max |d reward| | max |d ampl_val| (deg) | sweeps/step | "
          "worst row velocity error left by PGS (rad/s) | impulse-cap hits | limit rows | max prismatic |q| (m) | limit crossings")
    for fam, cs in cases(args.quick):
        print("\n== family %s (%d cases, %d steps)" % (fam, len(cs), sum(c[4] for c in cs)))
        ideal = []
        for c in cs:
            pol = actor if c[5] == "actor" else (lambda o, t, rs=np.random.RandomState(c[5]): G.f32_actions(rs))
            ideal.append(run(Env, bc, {}, c[0], c[1], c[2], c[3], c[4], pol)[0])
        for name, knobs in variants:
            t0 = time.time()
            agg = dict(dq=0.0, dobs=0.0, drew=0.0, damp=0.0, sweeps=0, steps=0, res=0.0, caps=0, lim=0, pris=0.0, cross=0)
            for c, ri in zip(cs, ideal):
                pol = actor if c[5] == "actor" else (lambda o, t, rs=np.random.RandomState(c[5]): G.f32_actions(rs))
                rf, w = run(Env, bc, knobs, c[0], c[1], c[2], c[3], c[4], pol)
                st = w.full[0].stats
                agg["dq"] = max(agg["dq"], np.abs(np.array(ri.st["qb"]) - np.array(rf.st["qb"])).max())
                agg["dobs"] = max(agg["dobs"], np.abs(np.array(ri.st["obs"]) - np.array(rf.st["obs"])).max(),
                                  np.abs(np.array(ri.ep["obs0"]) - np.array(rf.ep["obs0"])).max())
                agg["drew"] = max(agg["drew"], np.abs(np.array(ri.st["reward"]) - np.array(rf.st["reward"])).max())
                agg["damp"] = max(agg["damp"], np.abs(np.array(ri.st["ampl_val"]) - np.array(rf.st["ampl_val"])).max())
                agg["sweeps"] += st["sweeps"]; agg["steps"] += st["steps"]
                agg["res"] = max(agg["res"], st["max_residual"]); agg["caps"] += st["cap_hits"]; agg["lim"] += st["limit_rows"]
                agg["pris"] = max(agg["pris"], st["max_prismatic"]); agg["cross"] += w.limit_crossings
            print("%-112s | %.2e | %.2e | %.2e | %.2e | %5.1f | %.1e | %d | %d | %.1e | %d   (%.0f s)" % (
                name, agg["dq"], agg["dobs"], agg["drew"], agg["damp"], agg["sweeps"] / max(agg["steps"], 1), agg["res"],
                agg["caps"], agg["lim"], agg["pris"], agg["cross"], time.time() - t0))
            sys.stdout.flush()


if __name__ == "__main__":
    main()
