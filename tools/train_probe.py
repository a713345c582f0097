#!/usr/bin/env python
"""BASELINE config 5 end to end on one GPU: exo_b200.train_device (env -> policy + pink noise -> replay add -> TD7 iterations)
for a few population sizes. Prints per-episode mean returns (the learning signal) and whole-loop env-steps/s, learner included.

    python tools/train_probe.py [--replicas 1 64 1024] [--episodes 6]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import exo_b200  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--replicas", type=int, nargs="+", default=[1, 64, 1024])
    ap.add_argument("--episodes", type=int, default=6)
    ap.add_argument("--updates", type=float, default=1.0, help="TD7 iterations per episode time step")
    args = ap.parse_args()
    for rep in args.replicas:
        n = 8 * rep
        steps = int(n * 290 * args.episodes * 0.97)
        r = exo_b200.train_device(steps, replicas=rep, warmup=n * 290 + 1, device="cuda:0", seed=1, updates_per_episode_step=args.updates,
                                  learning_steps=6000000)
        print(json.dumps({"envs": n, "episodes": len(r["episode_return"]), "env_steps": r["env_steps"], "td7_iterations": r["train_iterations"],
                          "seconds": round(r["seconds"], 3), "env_steps_per_s": r["env_steps"] / r["seconds"],
                          "episode_mean_return": [round(x, 2) for x in r["episode_return"]],
                          "losses": {k: r["scalars"][k] for k in ("encoder_loss", "critic_loss", "actor_loss")}}))
        for k in ("agent", "policy", "env"):
            r[k].close()


if __name__ == "__main__":
    main()
