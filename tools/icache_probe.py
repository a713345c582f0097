#!/usr/bin/env python
"""Is part of exo_step_kernel's cold-start the INSTRUCTION fetch? The bench's flush also evicts the kernel's code from L2.
flush -> step(population A) -> step(population B): B's data is as cold as A's (different arrays, never touched since the
flush), but the code is back in L2 / the instruction caches. Also: flush -> a tiny 64-env step (code only) -> step(A).

    python tools/icache_probe.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import exo_b200  # noqa: E402

n = 65536
envs = []
for p in range(2):
    motion, seqs, seeds = bench.population(n, 0, 1)
    e = exo_b200.BatchedExoskeletonEnv(motion, seqs, [s + 1000003 * p for s in seeds], device="cuda:0", auto_reset=False, **bench.TRAIN)
    e.reset(); e.wait_prefetch()
    envs.append(e)
motion, seqs, seeds = bench.population(64, 0, 1)
tiny = exo_b200.BatchedExoskeletonEnv(motion, seqs, seeds, device="cuda:0", auto_reset=False, **bench.TRAIN)
tiny.reset(); tiny.wait_prefetch()
g = torch.Generator(device="cuda").manual_seed(0)
acts = torch.rand(4, n, 7, device="cuda", generator=g) * 2 - 1
tacts = acts[0, :64].contiguous()
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()


def run(seq, k=60):
    """seq: list of callables; events around each; returns mean us per position"""
    ev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in seq] for _ in range(k)]
    for t in range(k):
        flush.zero_()
        for j, f in enumerate(seq):
            ev[t][j][0].record(); f(t); ev[t][j][1].record()
    torch.cuda.synchronize()
    return [float(np.mean([ev[t][j][0].elapsed_time(ev[t][j][1]) for t in range(5, k)])) * 1e3 for j in range(len(seq))]


A = lambda t: envs[0].step(acts[t % 4], keep_idle_obs=False)
B = lambda t: envs[1].step(acts[(t + 1) % 4], keep_idle_obs=False)
T = lambda t: tiny.step(tacts, keep_idle_obs=False)
print("flush, A            : A %.2f us" % tuple(run([A])))
print("flush, A, B         : A %.2f  B %.2f us   (B: cold data, warm code, L2 partly holding A's results)" % tuple(run([A, B])))
print("flush, B, A         : B %.2f  A %.2f us" % tuple(run([B, A])))
print("flush, tiny, A      : tiny %.2f  A %.2f us   (A: cold data, code fetched by the 64-env launch)" % tuple(run([T, A])))
print("flush, tiny, tiny, A: tiny %.2f  tiny %.2f  A %.2f us" % tuple(run([T, T, A])))
