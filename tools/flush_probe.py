#!/usr/bin/env python
"""How the L2-flush method between timed launches changes the measured exo_step_kernel time.

    python tools/flush_probe.py [n_envs]

Variants: none (state stays L2 resident) | write (512 MiB overwrite: L2 left full of DIRTY lines, whose eviction
is charged to the next kernel) | write+read (overwrite, then read 512 MiB: L2 left full of CLEAN lines).
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
    actions = torch.rand(8, n, 7, device=dev) * 2 - 1
    wbuf = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    rbuf = torch.zeros(128 << 20, dtype=torch.float32, device=dev)
    env.reset()
    for mode in ("none", "write", "write+read", "read"):
        for rep in range(2):
            ts = []
            for t in range(60):
                if "write" in mode:
                    wbuf.zero_()
                if "read" in mode:
                    rbuf.sum()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                env.step(actions[t % 8], keep_idle_obs=False)
                b.record()
                ts.append((a, b))
            torch.cuda.synchronize()
            v = np.array([a.elapsed_time(b) for a, b in ts][10:]) * 1e3
            print("flush=%-10s rep %d  step kernel: mean %.2f us  median %.2f  min %.2f" % (mode, rep, v.mean(), np.median(v), v.min()))
        env.reset()
    # empty event pair = the floor the event method adds
    ts = []
    for t in range(50):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); b.record(); ts.append((a, b))
    torch.cuda.synchronize()
    print("empty event pair: %.2f us" % (np.mean([a.elapsed_time(b) for a, b in ts]) * 1e3))


if __name__ == "__main__":
    main()
