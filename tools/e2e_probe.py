#!/usr/bin/env python
"""Where exo_step_host's time goes: the PCIe copies alone (torch, pinned) next to the whole call."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import exo_b200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
dev = torch.device("cuda:0")
obs_d = torch.zeros(n, 80, device=dev); obs_h = torch.zeros(n, 80).pin_memory()
act_d = torch.zeros(n, 7, device=dev); act_h = torch.zeros(n, 7).pin_memory()


def wall(fn, k=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(k):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / k * 1e6


def d2h():
    obs_h.copy_(obs_d, non_blocking=True); torch.cuda.synchronize()


def h2d():
    act_d.copy_(act_h, non_blocking=True); torch.cuda.synchronize()


print("D2H obs %.1f MB: %.1f us   H2D actions %.1f MB: %.1f us" % (n * 320 / 1e6, wall(d2h), n * 28 / 1e6, wall(h2d)))
motion, seqs, seeds = bench.population(n, 0, 1)
env = exo_b200.BatchedExoskeletonEnv(motion, seqs, seeds, device="cuda:0", auto_reset=True, **bench.TRAIN)
env.reset()
acts = [(torch.rand(n, 7) * 2 - 1).pin_memory() for _ in range(4)]
i = [0]


def step():
    env.step_host(acts[i[0] % 4]); i[0] += 1


print("exo_step_host: %.1f us per call" % wall(step))
dact = acts[0].cuda()
print("env.step (device tensors) + sync: %.1f us" % wall(lambda: (env.step(dact), torch.cuda.synchronize())))
