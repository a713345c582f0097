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
    ap.add_argument("--dp", action="store_true", help="under torchrun: one rank per GPU, data-parallel learners over NCCL")
    args = ap.parse_args()
    rank, world, dev = 0, 1, "cuda:0"
    if args.dp:
        import torch
        import torch.distributed as dist
        rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
        torch.cuda.set_device(local)
        dev = "cuda:%d" % local
        dist.init_process_group("nccl", device_id=torch.device(dev))
    for rep in args.replicas:
        n = 8 * rep
        steps = int(n * 290 * args.episodes * 0.97)
        r = exo_b200.train_device(steps, replicas=rep, warmup=n * 290 + 1, device=dev, seed=1 + rank, updates_per_episode_step=args.updates,
                                  learning_steps=6000000, process_group=True if args.dp else None, init_seed=1)
        if args.dp:
            t = torch.tensor([r["seconds"]], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e = torch.tensor([float(r["env_steps"])], device=dev, dtype=torch.float64)
            dist.all_reduce(e)
            r["seconds"], r["env_steps"] = float(t), int(e)
        if rank == 0:
            print(json.dumps({"ranks": world, "envs": n * world, "episodes": len(r["episode_return"]), "env_steps": r["env_steps"], "td7_iterations": r["train_iterations"],
                          "seconds": round(r["seconds"], 3), "env_steps_per_s": r["env_steps"] / r["seconds"],
                          "episode_mean_return": [round(x, 2) for x in r["episode_return"]],
                              "losses": {k: r["scalars"][k] for k in ("encoder_loss", "critic_loss", "actor_loss")}}))
        for k in ("agent", "policy", "env"):
            r[k].close()
    if args.dp:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
