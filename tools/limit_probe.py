#!/usr/bin/env python
"""How often does a joint sit at / beyond a URDF limit in the bench population (decides which solver path the general
multibody step takes)?  python tools/limit_probe.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench, exo_b200
n = 65536
env = exo_b200.BatchedExoskeletonEnv(*bench.population(n, 0, 1), device="cuda:0", auto_reset=True, **bench.TRAIN)
g = torch.Generator(device="cuda").manual_seed(0)
acts = torch.rand(4, n, 7, device="cuda", generator=g) * 2 - 1
env.reset(); bench.dephase(env, acts); env.stats(reset=True)
lo = torch.tensor([-1.3962633609772, -0.69813168048859, -2.6441738605499, -0.034906584769487, -1.5184364318848], device="cuda")
hi = torch.tensor([1.3962633609772, 2.8187066316605, 0.78539800643921, 2.6179938726127, 1.3962633609772], device="cuda")
tot = viol = 0
per_joint = torch.zeros(5, device="cuda")
warps = 0
for t in range(100):
    env.step(acts[t % 4])
    q = env.get_field("q")
    v = (q <= lo) | (q >= hi)
    per_joint += v.float().sum(0)
    anyv = v.any(1)
    viol += int(anyv.sum()); tot += n
    warps += int(anyv.view(-1, 32).any(1).sum())
st = env.stats().cpu().numpy()
print("envs with a joint at/over a limit: %.4f of env-steps; warps containing one: %.4f; per joint %s; limit crossings stat %.0f of %.0f env-steps"
      % (viol / tot, warps / (tot / 32), (per_joint / tot).cpu().numpy(), st[4], st[3]))
