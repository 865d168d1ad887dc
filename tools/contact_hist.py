#!/usr/bin/env python
"""How many manifold points / solver rows / rolling-friction manifolds a tick carries on each workload, measured on the
(uncapped) CPU oracle under the benchmarked random policy. Decides the kernel's capacities.

    python tools/contact_hist.py [--envs 64] [--steps 40]
"""
import argparse
import json
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import oracle_lib as O  # noqa: E402


def actions(rng, name, n):
    if name == 'GolfJoints':
        return rng.uniform(-0.1, 0.1, (n, 6))
    if name == 'CubesGrasp':
        return np.concatenate([rng.uniform(0, 1, (n, 2)), rng.uniform(-np.pi, np.pi, (n, 6)), rng.uniform(0, 1, (n, 1))], 1)
    return rng.uniform(-0.1, 0.1, (n, 3))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--envs', type=int, default=64)
    ap.add_argument('--steps', type=int, default=40)
    ap.add_argument('--bias-down', type=float, default=0.0, help='shift of the z action (CubesPush / Golf): pushes the gripper into the table')
    ap.add_argument('names', nargs='*')
    a = ap.parse_args()
    thr = os.cpu_count() or 1
    for name in (a.names or ['BallBalance', 'TwoBallsBalance', 'CubesPush', 'CubesGrasp', 'Golf', 'GolfJoints']):
        n = a.envs
        ob = O.OracleBatch(name, n, seed=77)
        cuts = np.linspace(0, n, thr + 1).astype(int)
        rng = np.random.default_rng(5)
        with ThreadPoolExecutor(thr) as ex:
            list(ex.map(lambda i: ob.reset(lo=int(cuts[i]), hi=int(cuts[i + 1])), range(thr)))
            for k in range(a.steps):
                act = actions(rng, name, n)
                if a.bias_down and act.shape[1] == 3:
                    act[:, 2] -= a.bias_down
                list(ex.map(lambda i: ob.step(act, autoreset=True, lo=int(cuts[i]), hi=int(cuts[i + 1])), range(thr)))
        h = ob.tick_hist()
        c = h['contacts']; tot = c.sum()
        cum = np.cumsum(c) / tot
        out = dict(env=name, ticks=int(tot), contacts_max=int(np.nonzero(c)[0].max()), rows_max=int(np.nonzero(h['rows'])[0].max()),
                   rolling_max=int(np.nonzero(h['rolling'])[0].max()),
                   frac_ticks_contacts_gt={k: float(1 - cum[k]) for k in (8, 12, 16, 20, 24, 28, 32)},
                   rolling_hist=h['rolling'][:8].tolist())
        print(json.dumps(out), flush=True)


if __name__ == '__main__':
    main()
