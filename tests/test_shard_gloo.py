"""Multi-rank layout on CPU: shard ranges, the episode-statistics all-reduce over gloo (world_size 2), and the
layout-independence that makes weak scaling collective-free (a shard at index offset k reproduces envs k.. of the
whole batch bit for bit — checked with the oracle standing in for the device on each rank)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from urgym_b200 import shard


def test_shard_range_partitions_exactly():
    for total in (0, 1, 7, 8, 4096, 16385):
        for world in (1, 2, 3, 8):
            spans = [shard.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        shard.shard_range(8, 2, 2)


def test_reduce_stats_is_identity_without_a_process_group():
    st = dict(zip(shard.STAT_KEYS, range(6)))
    assert shard.reduce_stats(st) == st


def _free_port():
    s = socket.socket(); s.bind(('127.0.0.1', 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, total, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        import sys
        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        import oracle_lib as O
        r, w, _ = shard.rank_world()
        lo, hi = shard.shard_range(total, r, w)
        ob = O.OracleBatch('CubesGrasp', hi - lo, seed=17, offset=lo)       # stands in for the rank's device batch
        obs, nt = ob.reset()
        rng = np.random.default_rng(5)
        acts = rng.uniform(0, 1, (3, total, 9))
        episodes = ret = length = succ = steps = ticks = 0
        for k in range(3):
            out = ob.step(acts[k, lo:hi])
            steps += hi - lo; ticks += int(out['nticks'][:, 0].sum()); ret += float(out['reward'].sum())
        local = dict(zip(shard.STAT_KEYS, (episodes, ret, length, succ, steps, ticks)))
        glob = shard.reduce_stats(local)
        # gather the shards' observations to rank 0 for the layout check (test plumbing, not the data path)
        parts = [None] * w
        dist.all_gather_object(parts, (lo, hi, out['obs'], out['reward']))
        if r == 0:
            q.put((glob, parts))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_stats_and_layout_independence(built):
    import oracle_lib as O
    total, world = 6, 2
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    [p.start() for p in procs]
    glob, parts = q.get(timeout=240)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    # single-process run of the whole batch
    ob = O.OracleBatch('CubesGrasp', total, seed=17)
    ob.reset()
    rng = np.random.default_rng(5)
    acts = rng.uniform(0, 1, (3, total, 9))
    ticks = ret = 0
    for k in range(3):
        out = ob.step(acts[k]); ticks += int(out['nticks'][:, 0].sum()); ret += float(out['reward'].sum())
    assert glob['env_steps'] == 3 * total and glob['ticks'] == ticks
    assert glob['return_sum'] == pytest.approx(ret, abs=1e-9)
    for lo, hi, obs, rew in parts:
        np.testing.assert_array_equal(obs, out['obs'][lo:hi])
        np.testing.assert_array_equal(rew, out['reward'][lo:hi])
