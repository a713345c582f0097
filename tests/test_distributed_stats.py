"""CPU, world_size 2, gloo: the N>1 host path — shard the population, reduce the episode-statistics vector."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import REPO


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_total, out):
    sys.path.insert(0, REPO)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from exo_b200 import sharding
    motion, seqs, seeds = sharding.benchmark_population(n_total, rank, world)
    lo, hi = sharding.shard_range(n_total, rank, world)
    stats = torch.zeros(16, dtype=torch.float64)
    stats[0] = hi - lo                       # pretend every env finished one episode
    stats[1] = float(seeds.astype(np.float64).sum())
    stats[3] = (hi - lo) * 10
    sharding.reduce_stats(stats)
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)    # max-over-ranks timing rule used by bench.py
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        torch.save(dict(stats=stats, tmax=t, n_local=hi - lo, motions=np.bincount(motion, minlength=8)), out)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_and_stats_reduction(tmp_path):
    out = str(tmp_path / "r0.pt")
    n_total = 1001
    mp.spawn(_worker, args=(2, _free_port(), n_total, out), nprocs=2, join=True)
    r = torch.load(out, weights_only=False)
    assert r["stats"][0].item() == n_total
    assert r["stats"][1].item() == n_total * (n_total - 1) / 2          # every global env id counted exactly once
    assert r["stats"][3].item() == n_total * 10
    assert r["tmax"].item() == 2.0
    assert r["n_local"] == 501 and r["motions"].sum() == 501
