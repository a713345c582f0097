#!/usr/bin/env python
"""ncu target: exo_step_kernel launched the way bench.py times it - each launch preceded by the L2 flush (zeroing a 512 MiB
buffer), which leaves L2 full of DIRTY lines that the step kernel's own allocations push out to DRAM. Profile with
`--cache-control none` (ncu's default invalidates L2 before a launch and shows a clean-cache kernel instead):

    ncu --cache-control none --clock-control none -k regex:exo_step_kernel -s 5 -c 12 \
        --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum python tools/ncu_target_dirty_l2.py [write|read]

argument `read`: the flush is followed by a read pass over the same buffer (L2 cold but clean)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench, exo_b200
n = 65536
mode = sys.argv[1] if len(sys.argv) > 1 else "write"
env = exo_b200.BatchedExoskeletonEnv(*bench.population(n, 0, 1), device="cuda:0", auto_reset=False, **bench.TRAIN)
g = torch.Generator(device="cuda").manual_seed(0)
acts = torch.rand(4, n, 7, device="cuda", generator=g) * 2 - 1
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
env.reset(); env.wait_prefetch(); torch.cuda.synchronize()
for t in range(20):
    flush.zero_()
    if mode == "read":
        flush.view(torch.int32).sum()
    env.step(acts[t % 4], keep_idle_obs=False)
torch.cuda.synchronize()
print("ok")
