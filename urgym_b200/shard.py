"""Multi-GPU layout: envs are independent, so each rank owns a contiguous shard and the step path has no collective.

The only exchange is the optional sum of the six episode counters (episodes, return sum, length sum, successes, env
steps, physics ticks) — 48 bytes per rank, one all-reduce (NCCL on GPU tensors; gloo in the CPU tests). SURVEY.md §8e.
"""
import os

STAT_KEYS = ('episodes', 'return_sum', 'length_sum', 'successes', 'env_steps', 'ticks')


def shard_range(total_envs, rank, world_size):
    """Contiguous global index range [lo, hi) of `rank`; remainders go to the lowest ranks."""
    if not (0 <= rank < world_size) or total_envs < 0:
        raise ValueError('bad shard request')
    base, rem = divmod(total_envs, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def rank_world():
    return int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1')), int(os.environ.get('LOCAL_RANK', '0'))


def reduce_stats(stats, device=None):
    """Sum a stats dict over all ranks of the default process group (no-op when torch.distributed is not initialised)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return dict(stats)
    dev = device if (device is not None and dist.get_backend() == 'nccl') else 'cpu'
    t = torch.tensor([float(stats[k]) for k in STAT_KEYS], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return dict(zip(STAT_KEYS, t.tolist()))


def make_sharded(env_id, total_envs, seed=0, **kwargs):
    """One URGymVectorEnv per rank over its shard of `total_envs` global envs (results do not depend on the layout:
    the RNG is keyed on the global env index)."""
    from .envs import URGymVectorEnv
    rank, world, local = rank_world()
    lo, hi = shard_range(total_envs, rank, world)
    return URGymVectorEnv(env_id, hi - lo, device=local, seed=seed, env_index_offset=lo, **kwargs)
