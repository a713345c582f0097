#!/usr/bin/env python
"""Data-parallel TD7 learners over NCCL: one rank per GPU, each with its own batch, gradients all-reduced before Adam.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/td7_dp_probe.py

Prints ms per iteration (CUDA events, max over ranks), whole-job transitions/s, and checks that the ranks' networks stay
bit-identical (a checksum all-gathered at the end)."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import exo_b200  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    hp = exo_b200.Hyperparameters()
    t = exo_b200.DeviceTD7(80, 7, 1, hp=hp, env_num=15, device=dev, replay_buffer=False, init_seed=1, process_group=True)
    B = t.batch_rows
    g = torch.Generator(device=dev).manual_seed(100 + rank)           # every rank its own data
    batches = [[torch.randn(B, 80, device=dev, generator=g) * 0.6, (torch.randn(B, 7, device=dev, generator=g) * 0.5).clamp(-1, 1),
                torch.randn(B, 80, device=dev, generator=g) * 0.6, torch.randn(B, 1, device=dev, generator=g) * 3 + 1,
                (torch.rand(B, 1, device=dev, generator=g) > 0.1).float()] for _ in range(8)]
    for i in range(20):
        t.train_on_batch(*batches[i % 8])
    iters = 200
    dist.barrier()
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(iters):
        t.train_on_batch(*batches[i % 8])
    b.record()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([a.elapsed_time(b) / iters], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    chk = torch.stack([torch.cat([v.double().reshape(-1) for v in t.state_dict(n).values()]).sum() for n in ("actor", "critic", "encoder")]).to(dev)
    allc = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(allc, chk)
    same = all(torch.equal(allc[0], c) for c in allc)
    if rank == 0:
        print("ranks %d: %.3f ms per iteration (max over ranks), %.3e transitions/s whole job, networks identical across ranks: %s"
              % (world, float(ms), world * B / (float(ms) * 1e-3), same))
    t.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
