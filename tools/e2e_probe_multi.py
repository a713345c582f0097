#!/usr/bin/env python
"""Ceiling of the host-buffer path on N GPUs of one box: bare concurrent pinned device->host copies of one step's
observations (21 MB per rank) from every rank at once, next to exo_step_host itself.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/e2e_probe_multi.py [--pin 0|1] [--chunks K]

Prints, per N: GB/s per rank and aggregate of the bare copy (what the PCIe links / host memory give N ranks at once), the
env-steps/s that ceiling allows (n x 80 x 4 B of observations per step dominate the traffic), and the measured exo_step_host
rate as a fraction of it."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import exo_b200  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--pin", type=int, default=1)
ap.add_argument("--chunks", type=int, default=0)
ap.add_argument("--envs", type=int, default=65536)
args = ap.parse_args()
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
pin = exo_b200.hostpin.pin_to_gpu(lr) if args.pin else {"pinned": False}
if args.chunks:
    os.environ["EXO_HOST_CHUNKS"] = str(args.chunks)
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
n = args.envs


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)


def maxr(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


obs_d = torch.zeros(n, 80, device=dev)
obs_h = torch.zeros(n, 80).pin_memory()
reps = 40
for _ in range(5):
    obs_h.copy_(obs_d, non_blocking=True)
barrier()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(reps):
    obs_h.copy_(obs_d, non_blocking=True)
b.record()
barrier()
ms = maxr(a.elapsed_time(b)) / reps
gbs = n * 320 / (ms * 1e-3) / 1e9
env = exo_b200.BatchedExoskeletonEnv(*bench.population(n, rank, world), device=dev, auto_reset=True, **bench.TRAIN)
env.reset(); env.wait_prefetch(); torch.cuda.synchronize()
acts = [(torch.rand(n, 7) * 2 - 1).pin_memory() for _ in range(4)]
for t in range(5):
    env.step_host(acts[t % 4])
barrier()
a.record()
for t in range(reps):
    env.step_host(acts[t % 4])
b.record()
barrier()
ms2 = maxr(a.elapsed_time(b)) / reps
if rank == 0:
    ceil = world * n / (ms * 1e-3)
    got = world * n / (ms2 * 1e-3)
    print("N=%d pin=%d chunks=%s (%s): bare D2H %.1f GB/s per rank, %.1f GB/s aggregate -> ceiling %.3e env-steps/s; exo_step_host %.3e "
          "env-steps/s = %.2f of it (%.1f us per call)" % (world, args.pin, args.chunks or "default", pin, gbs, gbs * world, ceil, got, got / ceil, ms2 * 1e3))
if world > 1:
    dist.destroy_process_group()
