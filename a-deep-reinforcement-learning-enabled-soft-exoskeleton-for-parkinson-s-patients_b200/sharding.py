"""Multi-GPU layout of the env population: independent envs, one contiguous shard per rank, no per-step
collective. The only exchange is a sum all-reduce of the 16-entry episode-statistics vector (NCCL on GPUs,
gloo in the CPU tests)."""
import numpy as np

TYPES15 = [[(m >> 3) & 1, (m >> 2) & 1, (m >> 1) & 1, m & 1, 0, 0, 0] for m in range(1, 16)]


def shard_range(n_total, rank, world):
    """[lo, hi) of the global env ids owned by `rank` (sizes differ by at most one)."""
    base, extra = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


MOTION_BLOCK = 1024      # envs per contiguous same-motion block (a multiple of the 32-env warp and of the 64-env step block)


def benchmark_population(n_total, rank=0, world=1, n_motions=8):
    """BASELINE configs 3/4 population for this rank's shard. Everything is a function of the GLOBAL env id g, so global
    env g is the same env (motion, tremor type, seed -> every RNG-derived episode parameter, every trajectory) whatever the
    world size: motion = (g // 1024) mod 8 (contiguous 1024-env blocks per reference motion, so a warp shares its motion row
    and episode phase and every shard holds an even mix of episode lengths), tremor type = the 15 non-zero [a,b,c,d]
    patterns round-robin over g, seed = g."""
    lo, hi = shard_range(n_total, rank, world)
    gid = np.arange(lo, hi)
    motion = ((gid // MOTION_BLOCK) % n_motions).astype(np.int32)
    seqs = np.asarray(TYPES15, dtype=np.uint8)[gid % 15]
    return motion, seqs, gid.astype(np.uint32)


def reduce_stats(stats, group=None):
    """Sum the per-rank statistics vector over all ranks in place (no-op without an initialised process group)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def summarize(stats):
    s = [float(x) for x in stats]
    ep = max(s[0], 1.0)
    return dict(episodes=s[0], mean_return=s[1] / ep, mean_length=s[2] / ep, env_steps=s[3], limit_crossings=s[4],
                frac_steps_with_arm_reduction=s[5] / max(s[3], 1.0),
                mean_negative_amp_metric=s[6] / max(s[7], 1.0),
                # the training script's console figures (Exoskeleton_agent_train.py:232-262), over all envs of the shard
                tremor_occurrence_pct=100.0 * s[9] / max(s[8] + s[9], 1.0),
                tremor_suppression_avg_pct=s[10] / max(s[11], 1.0),
                tremor_ampl_suppression_avg_pct=s[12] / max(s[13], 1.0),
                tremor_ampl_total_occurrence_pct=100.0 * s[7] / max(s[3], 1.0))
