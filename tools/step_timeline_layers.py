#!/usr/bin/env python
"""exo_step_kernel timeline by dispatch layer (blockIdx / n_sm): absolute time of each mark since the first warp's entry,
for one EXO_STEP_STAGGER setting (the knob is read when the handle is created: one process per setting).

    EXO_STEP_STAGGER=1470 python tools/step_timeline_layers.py [n_envs]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import exo_b200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
motion, seqs, seeds = bench.population(n, 0, 1)
env = exo_b200.BatchedExoskeletonEnv(motion, seqs, seeds, device="cuda:0", auto_reset=True, **bench.TRAIN)
env.reset(); env.wait_prefetch(); torch.cuda.synchronize()
g = torch.Generator(device="cuda").manual_seed(0)
acts = torch.rand(4, n, 7, device="cuda", generator=g) * 2 - 1
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
for t in range(5):
    env.step(acts[t % 4], keep_idle_obs=False)
nw = (n + 31) // 32
buf = torch.zeros(nw, 8, dtype=torch.int64, device="cuda")
evs = []
for rep in range(6):
    flush.zero_()
    env.lib.exo_debug_timeline(env.h, C.c_void_p(buf.data_ptr()))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); env.step(acts[rep % 4], keep_idle_obs=False); b.record()
    torch.cuda.synchronize()
    env.lib.exo_debug_timeline(env.h, None)
    evs.append(a.elapsed_time(b) * 1e3)
    tl = buf.cpu().numpy().copy()
pl = []                                                             # plain launches: the event time without instrumentation
for rep in range(20):
    flush.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); env.step(acts[rep % 4], keep_idle_obs=False); b.record()
    torch.cuda.synchronize()
    pl.append(a.elapsed_time(b) * 1e3)
ghz = 1.965
t_entry = (tl[:, 0] - tl[:, 0].min()) / 1e3
marks = t_entry[:, None] + (tl[:, 2:8] - tl[:, 2:3]) / ghz / 1e3       # absolute us
layer = (np.arange(nw) // 2) // 148
print("stagger %s: events %.1f us instrumented, %.1f us plain (mean of 20); last exit %.2f us" % (
    os.environ.get("EXO_STEP_STAGGER", "default"), np.median(evs), np.mean(pl), marks[:, 5].max()))
print("layer  warps | median absolute us: entry, loads issued, data landed, actuators done, reward done, obs written | max exit")
for l in np.unique(layer):
    m = marks[layer == l]
    print("%5d %6d | %s | %.2f" % (l, len(m), "  ".join("%5.2f" % v for v in np.median(m, axis=0)), m[:, 5].max()))
