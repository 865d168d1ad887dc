#!/usr/bin/env python
"""Record the CPU oracle's side of tests/test_gpu_parity.py::test_episode_statistics_over_1e4_episodes:
>= 1e4 finished episodes per env under the test's seeded random policy (mean / variance of return, length, success).
The oracle does not depend on the GPU, so its 40 M ticks are spent here once instead of on the GPU box at every run.
Regenerate whenever oracle/ changes: python tests/golden/make_episode_stats.py  (a few minutes on 8 cores)."""
import json
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as O  # noqa: E402

CONFIGS = [('BallBalance', 4096, 56), ('CubesPush', 1024, 500), ('CubesGrasp', 1024, 200)]   # (env, envs, steps): the test's table
SEED_ENV, SEED_ACTIONS = 41, 11


def actions(rng, name, n):   # tests/test_gpu_parity.py::_actions
    if name == 'CubesGrasp':
        return np.concatenate([rng.uniform(0, 1, (n, 2)), rng.uniform(-np.pi, np.pi, (n, 6)), rng.uniform(0, 1, (n, 1))], 1).astype(np.float32).astype(np.float64)
    return rng.uniform(-0.1, 0.1, (n, 3)).astype(np.float32).astype(np.float64)


def main():
    O.build()
    out = {}
    for name, n, steps in CONFIGS:
        ob = O.OracleBatch(name, n, seed=SEED_ENV)
        rng = np.random.default_rng(SEED_ACTIONS)
        nthr = min(32, os.cpu_count() or 1)
        cuts = np.linspace(0, n, nthr + 1).astype(int)
        ret, length, eps = np.zeros(n), np.zeros(n, int), []
        with ThreadPoolExecutor(nthr) as ex:
            list(ex.map(lambda i: ob.reset(lo=int(cuts[i]), hi=int(cuts[i + 1])), range(nthr)))
            for k in range(steps):
                a = actions(rng, name, n); u = rng.uniform(0, 1, n).astype(np.float32).astype(np.float64)
                res = list(ex.map(lambda i: ob.step(a, u=u, autoreset=True, lo=int(cuts[i]), hi=int(cuts[i + 1])), range(nthr)))
                rew = np.zeros(n); done = np.zeros(n, bool); suc = np.zeros(n, bool)
                for i, r in enumerate(res):
                    sl = slice(int(cuts[i]), int(cuts[i + 1]))
                    rew[sl] = r['reward'][sl]; done[sl] = r['terminated'][sl] | r['truncated'][sl]; suc[sl] = r['success'][sl]
                ret += rew; length += 1
                for i in np.nonzero(done)[0]:
                    eps.append((ret[i], length[i], suc[i]))
                ret[done] = 0; length[done] = 0
        e = np.array(eps, float)
        out[name] = dict(envs=n, steps=steps, seed_env=SEED_ENV, seed_actions=SEED_ACTIONS, episodes=len(e),
                         mean=dict(zip(('return', 'length', 'success'), e.mean(0).tolist())), var=dict(zip(('return', 'length', 'success'), e.var(0).tolist())))
        print(name, json.dumps(out[name]), flush=True)
    with open(os.path.join(HERE, 'episode_stats_oracle.json'), 'w') as f:
        json.dump(out, f, indent=1)


if __name__ == '__main__':
    main()
