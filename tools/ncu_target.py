#!/usr/bin/env python
"""Short program for ncu captures of the env kernels at BASELINE config 3's size (65 536 envs, fp32):

    python tools/ncu_target.py [n_envs] [steps] [motor_model]

exo_seed's generation pass is the first exo_gen_kernel launch (every env), reset() the first exo_begin_kernel launch, then
`steps` exo_step_kernel launches from the common start (all warps live, no episode ends)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import exo_b200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
mm = sys.argv[3] if len(sys.argv) > 3 else "ideal"
motion, seqs, seeds = bench.population(n, 0, 1)
env = exo_b200.BatchedExoskeletonEnv(motion, seqs, seeds, device="cuda:0", auto_reset=True, motor_model=mm, **bench.TRAIN)
env.reset()
g = torch.Generator(device="cuda").manual_seed(0)
a = torch.rand(n, 7, device="cuda", generator=g) * 2 - 1
for t in range(steps):
    env.step(a, keep_idle_obs=False)
torch.cuda.synchronize()
print("ok", env.launch_count)
