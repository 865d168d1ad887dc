#!/usr/bin/env python
"""Replay the reference's shipped SB3 actors (trained on pybullet) on the CPU oracle and compare task statistics.

The checkpoints under /root/reference/policies_final/*/SAC_policy.zip hold actor weights (MLP obs -> 256 -> 256 ->
act, ReLU, tanh-squashed mean) and, in their `data` JSON, the Monitor buffers of the last 100 training episodes on
pybullet (return, length, success). If the oracle's physics behaves like Bullet's, a policy trained there must also
solve the task here. This is the only end-to-end pin of the rigid-body arithmetic available without pybullet
(SURVEY.md §8c "weak pins" 3). Nothing from the checkpoints is copied into the repo; run where /root/reference exists:

    python tools/replay_actors.py [--episodes 200] [--out profiles/r01_actor_replay.json]
"""
import argparse
import io
import json
import os
import sys
import zipfile
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import oracle_lib as O  # noqa: E402

REF = os.environ.get('URGYM_REFERENCE', '/root/reference')
CASES = {'CubesPush': 'policies_final/cubes_push_v0/SAC_policy.zip', 'Golf': 'policies_final/golf_v0/SAC_policy.zip',
         'GolfJoints': 'policies_final/golf_joints_v0/SAC_policy.zip'}


def load_actor(path):
    z = zipfile.ZipFile(path)
    sd = torch.load(io.BytesIO(z.read('policy.pth')), map_location='cpu', weights_only=True)
    data = json.loads(z.read('data'))
    g = lambda k: sd[k].double().numpy()
    W0, b0, W1, b1 = g('actor.latent_pi.0.weight'), g('actor.latent_pi.0.bias'), g('actor.latent_pi.2.weight'), g('actor.latent_pi.2.bias')
    sde = 'actor.mu.0.weight' in sd          # gSDE actors: mu = Linear + Hardtanh(-2, 2)
    Wm, bm = (g('actor.mu.0.weight'), g('actor.mu.0.bias')) if sde else (g('actor.mu.weight'), g('actor.mu.bias'))

    def act(obs):  # deterministic SAC action, rescaled to the env's [-0.1, 0.1] box
        x = np.asarray(obs, np.float32).astype(np.float64)
        h = np.maximum(x @ W0.T + b0, 0); h = np.maximum(h @ W1.T + b1, 0)
        m = h @ Wm.T + bm
        if sde:
            m = np.clip(m, -2.0, 2.0)
        return 0.1 * np.tanh(m)
    hist = dict(ep_info=data.get('ep_info_buffer'), ep_success=data.get('ep_success_buffer'), num_timesteps=data.get('num_timesteps'))
    return act, hist


def decode_buffer(b):
    """SB3 serialises deques through cloudpickle+base64; the readable part we need is absent, so unpickle."""
    import base64
    import pickle
    if not b or ':serialized:' not in b:
        return None
    try:
        return list(pickle.loads(base64.b64decode(b[':serialized:'])))
    except Exception as e:  # cloudpickle-only payloads
        return None


def rollout(name, act, episodes, seed, threads):
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
    return dict(episodes=n, mean_return=float(ret.mean()), mean_length=float(length.mean()), success_rate=float(succ.mean()))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--episodes', type=int, default=200)
    ap.add_argument('--out', default='')
    args = ap.parse_args()
    O.build()
    results = {}
    for name, rel in CASES.items():
        act, hist = load_actor(os.path.join(REF, rel))
        info, suc = decode_buffer(hist['ep_info']), decode_buffer(hist['ep_success'])
        trained = None
        if info:
            trained = dict(episodes=len(info), mean_return=float(np.mean([e['r'] for e in info])), mean_length=float(np.mean([e['l'] for e in info])),
                           success_rate=float(np.mean(suc)) if suc else None)
        r = rollout(name, act, args.episodes, seed=123, threads=min(16, os.cpu_count() or 1))
        results[name] = dict(oracle_replay=r, pybullet_training_last_100=trained, checkpoint=rel, num_timesteps=hist['num_timesteps'])
        print(name, json.dumps(results[name]), flush=True)
    if args.out:
        with open(args.out, 'w') as f:
            json.dump(results, f, indent=1)


if __name__ == '__main__':
    main()
