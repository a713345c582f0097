#!/usr/bin/env python
"""ncu target: one exo_step_kernel launch in the de-phased steady state (~232 lanes start a new episode inside it).
    python tools/ncu_target_dephased.py   ->  launch index 347 + 3 of exo_step_kernel is a steady-state one"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench, exo_b200
n = 65536
os.environ.setdefault("EXO_DEBUG_NO_GEN", "1")       # keep the launch list to step / begin kernels
mm = sys.argv[1] if len(sys.argv) > 1 else "ideal"
env = exo_b200.BatchedExoskeletonEnv(*bench.population(n, 0, 1), device="cuda:0", auto_reset=True, motor_model=mm, **bench.TRAIN)
g = torch.Generator(device="cuda").manual_seed(0)
acts = torch.rand(4, n, 7, device="cuda", generator=g) * 2 - 1
env.reset(); bench.dephase(env, acts)
for t in range(6):
    env.step(acts[t % 4], keep_idle_obs=False)
torch.cuda.synchronize()
print("ok")
