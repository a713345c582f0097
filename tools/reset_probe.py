#!/usr/bin/env python
"""Latency of exo_reset_kernel as a function of how many envs it resets in one launch.

    python tools/reset_probe.py [n_envs]

At 65 536 envs about 190 episodes finish per step; the step that follows waits for their reset. The kernel walks one
env per warp, so for small work lists its duration is one warp's serial chain, not a throughput figure. Masked
`reset(mask)` launches (L2 flushed before each) are timed with CUDA events for k = 0 ... n reset envs. The masked path
strides over all n envs and tests the mask (the auto-reset path walks a compact work list instead), so k = 0 is the
floor to subtract.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import exo_b200  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    dev = torch.device("cuda", 0)
    motion, seqs, seeds = bench.population(n, 0, 1)
    env = exo_b200.BatchedExoskeletonEnv(motion, seqs, seeds, device=dev, auto_reset=False, **bench.TRAIN)
    env.reset()
    wbuf = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    g = torch.Generator(device="cpu").manual_seed(0)
    for k in (0, 1, 32, 148, 190, 592, 2368, 16384, n):
        ts = []
        for rep in range(24):
            mask = torch.zeros(n, dtype=torch.uint8)
            if k:
                mask[torch.randperm(n, generator=g)[:k]] = 1
            mask = mask.to(dev)
            wbuf.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            env.reset(mask)
            env.wait_prefetch()          # the generation pass runs on the handle's side stream
            b.record()
            ts.append((a, b))
        torch.cuda.synchronize()
        v = np.array([a.elapsed_time(b) for a, b in ts][4:]) * 1e3
        print("reset of %6d envs: mean %8.2f us  median %8.2f  min %8.2f  (%.2f us per env)"
              % (k, v.mean(), np.median(v), v.min(), np.median(v) / max(k, 1)))


if __name__ == "__main__":
    main()
