#!/usr/bin/env python
"""Phase timeline of one exo_step_kernel launch (exo_debug_timeline): where a 65 536-env step spends its ~18 us.

    python tools/step_timeline.py [n_envs]

Per warp: entry time (globaltimer), SM, clock64 at six marks. Prints, over all warps, the distribution of the entry offsets
(launch ramp) and of each phase's duration, and the per-SM span from first entry to last exit."""
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
res = []
for rep in range(5):
    flush.zero_()
    env.lib.exo_debug_timeline(env.h, C.c_void_p(buf.data_ptr()))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    env.step(acts[rep % 4], keep_idle_obs=False)
    b.record()
    torch.cuda.synchronize()
    env.lib.exo_debug_timeline(env.h, None)
    res.append((a.elapsed_time(b) * 1e3, buf.cpu().numpy().copy()))
ev_us, tl = res[-1]
ghz = 1.965
t_entry = (tl[:, 0] - tl[:, 0].min()) / 1e3                  # us
marks = (tl[:, 2:8] - tl[:, 2:3]) / ghz / 1e3                 # us since the warp's own entry
names = ["entry", "loads issued", "data landed", "actuators done", "reward done", "obs written"]
print("events: %.1f us for the launch; warps %d" % (ev_us, nw))
print("entry offset (launch ramp): p0 %.2f  p50 %.2f  p90 %.2f  p100 %.2f us" % tuple(np.percentile(t_entry, [0, 50, 90, 100])))
for k in range(1, 6):
    d = marks[:, k] - marks[:, k - 1]
    print("%-16s -> %-16s p10 %.2f  p50 %.2f  p90 %.2f  max %.2f us" % ((names[k - 1], names[k]) + tuple(np.percentile(d, [10, 50, 90, 100]))))
life = marks[:, 5]
print("warp lifetime: p10 %.2f p50 %.2f p90 %.2f max %.2f us" % tuple(np.percentile(life, [10, 50, 90, 100])))
end = t_entry + life
print("last warp exit at %.2f us after the first entry" % end.max())
sm = tl[:, 1]
spans = [(end[sm == s].max() - t_entry[sm == s].min(), (sm == s).sum()) for s in np.unique(sm)]
print("per-SM span (first entry -> last exit): min %.2f  median %.2f  max %.2f us; warps per SM min %d max %d" % (
    min(s for s, _ in spans), float(np.median([s for s, _ in spans])), max(s for s, _ in spans), min(c for _, c in spans), max(c for _, c in spans)))
