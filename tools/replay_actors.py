#!/usr/bin/env python
"""Replay the reference's shipped SB3 actors (trained on pybullet) and compare task statistics with the checkpoints'.

The actors (tests/golden/actors.npz) are MLPs trained against pybullet; tests/golden/actors.json holds the Monitor
statistics of their last 100 training episodes there. If the physics behaves like Bullet's, a policy trained there must
also solve the task here: the only end-to-end pin of the rigid-body arithmetic available without pybullet.

    python tools/replay_actors.py [--episodes 1000] [--backend oracle|cuda] [--set name=value ...] [--out file.json] [env ...]

--backend oracle: the fp64 CPU restatement, all host threads (--set overrides recalled parameters, oracle/bullet_params.h).
--backend cuda:   the CUDA library; the MLP runs in torch on the device and drives the batch with no host synchronisation
                  on the step path (one `alive.any()` every 16 steps to stop early). cuda64 = the fp64 parity build.
"""
import argparse
import json
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import actors as A  # noqa: E402

ENVS = ['TwoBallsBalance', 'CubesPush', 'Golf', 'GolfJoints']


def rollout_oracle(name, episodes, seed=123, threads=None):
    import oracle_lib as O
    act = A.numpy_actor(name)
    threads = threads or min(32, os.cpu_count() or 1)
    n = episodes
    ob = O.OracleBatch(name, n, seed=seed)
    cuts = np.linspace(0, n, threads + 1).astype(int)
    with ThreadPoolExecutor(threads) as ex:
        res = list(ex.map(lambda i: ob.reset(lo=int(cuts[i]), hi=int(cuts[i + 1])), range(threads)))
        obs = np.zeros((n, ob.od))
        for i, (o, _) in enumerate(res):
            obs[cuts[i]:cuts[i + 1]] = o[cuts[i]:cuts[i + 1]]
        alive = np.ones(n, bool); ret = np.zeros(n); length = np.zeros(n, int); succ = np.zeros(n, bool)
        for k in range(ob.max_steps):
            a = act(obs)
            out = list(ex.map(lambda i: ob.step(a, lo=int(cuts[i]), hi=int(cuts[i + 1])), range(threads)))
            for i, r in enumerate(out):
                sl = slice(int(cuts[i]), int(cuts[i + 1]))
                live = alive[sl]
                obs[sl] = np.where(live[:, None], r['obs'][sl], obs[sl])
                ret[sl] += np.where(live, r['reward'][sl], 0); length[sl] += live
                done = (r['terminated'][sl] | r['truncated'][sl]) & live
                succ[sl] |= r['success'][sl] & live
                alive[sl] &= ~done
            if not alive.any():
                break
    return summarize(ret, length, succ)


def rollout_cuda(name, episodes, seed=123, precision='f32'):
    """First episode of `episodes` envs; everything stays on the device until the final statistics."""
    import torch
    from urgym_b200.batch import EnvBatch
    dev = torch.device('cuda', 0)
    act = A.torch_actor(name, dev)
    b = EnvBatch(name, episodes, seed=seed, precision=precision)
    obs = b.reset().clone()
    alive = torch.ones(episodes, dtype=torch.bool, device=dev)
    ret = torch.zeros(episodes, dtype=torch.float64, device=dev); length = torch.zeros(episodes, dtype=torch.int64, device=dev)
    succ = torch.zeros(episodes, dtype=torch.bool, device=dev)
    for k in range(b.max_episode_steps):
        if k % 16 == 15 and not bool(alive.any()):      # the only host synchronisation: once every 16 steps, to stop early
            break
        a = act(obs)
        o, r, term, trunc, su = b.step(a, autoreset=False)
        obs = torch.where(alive[:, None], o, obs)
        ret += torch.where(alive, r.double(), torch.zeros_like(ret)); length += alive
        succ |= su.bool() & alive
        alive &= ~(term.bool() | trunc.bool())
    out = summarize(ret.cpu().numpy(), length.cpu().numpy(), succ.cpu().numpy())
    st = b.stats()
    out['capacity_overflow_ticks'] = st['capacity_overflow_ticks']; out['ticks'] = st['ticks']
    b.close()
    return out


def summarize(ret, length, succ):
    n = len(ret)
    return dict(episodes=int(n), mean_return=float(ret.mean()), mean_length=float(length.mean()), success_rate=float(succ.mean()),
                se_return=float(ret.std() / np.sqrt(n)), se_length=float(length.std() / np.sqrt(n)), se_success=float(succ.std() / np.sqrt(n)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--episodes', type=int, default=1000)
    ap.add_argument('--backend', default='oracle', choices=['oracle', 'cuda', 'cuda64'])
    ap.add_argument('--set', action='append', default=[], help='oracle parameter override name=value (bullet_params.h: BulletParams)')
    ap.add_argument('--out', default='')
    ap.add_argument('names', nargs='*')
    args = ap.parse_args()
    if args.backend == 'oracle':
        import oracle_lib as O
        O.build()
        for kv in args.set:
            k, v = kv.split('=')
            O.set_param(k, float(v))
    results = {}
    for name in (args.names or ENVS):
        if args.backend == 'oracle':
            r = rollout_oracle(name, args.episodes)
        else:
            r = rollout_cuda(name, args.episodes, precision='f64' if args.backend == 'cuda64' else 'f32')
        results[name] = dict(replay=r, backend=args.backend, overrides=args.set, pybullet_training_last_100=A.META[name]['pybullet_training_last_100'],
                             checkpoint=A.META[name]['checkpoint'])
        print(name, json.dumps(results[name]), flush=True)
    if args.out:
        with open(args.out, 'w') as f:
            json.dump(results, f, indent=1)


if __name__ == '__main__':
    main()
