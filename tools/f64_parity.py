#!/usr/bin/env python
"""GPU-vs-oracle agreement of the integer outputs (settle trip counts, flags, contact-pair sets), product build (fp32)
and parity build (fp64) side by side, re-synchronised every gym step. Every mismatch is printed with the oracle's
decision margin: |summed joint motion - 5e-3| at the tick where one side stopped, or the smallest inequality slack of the
last tick's narrowphase.

    python tools/f64_parity.py [--envs 32] [--steps 50] [--out gpurun_out/f64_parity.jsonl] [env ...]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import oracle_lib as O  # noqa: E402
from urgym_b200.batch import EnvBatch  # noqa: E402

ENVS = ['BallBalance', 'TwoBallsBalance', 'CubesPush', 'CubesGrasp', 'Golf', 'CubesPushDiff', 'GolfJoints']


def actions(rng, name, n):
    if name == 'GolfJoints':
        return rng.uniform(-0.1, 0.1, (n, 6))
    if name == 'CubesGrasp':  # the benchmarked distribution: env_cubes_grasp_v0.py:74
        return np.concatenate([rng.uniform(0, 1, (n, 2)), rng.uniform(-np.pi, np.pi, (n, 6)), rng.uniform(0, 1, (n, 1))], 1)
    return rng.uniform(-0.1, 0.1, (n, 3))


def run(name, precision, n, steps, seed=7):
    ob, gb = O.OracleBatch(name, n, seed=seed), EnvBatch(name, n, seed=seed, precision=precision)
    ob.reset(); gb.reset()
    rng = np.random.default_rng(2)
    npdt = np.float32 if precision == 'f32' else np.float64
    tot = same = flags_ok = con_ok = 0
    tick_margins, con_margins, obs_err, tdiff = [], [], [], []
    for k in range(steps):
        s = ob.get_state().astype(npdt)
        gb.set_state(torch.tensor(s)); ob.set_state(s.astype(np.float64))
        a = actions(rng, name, n).astype(np.float32).astype(np.float64)
        u = rng.uniform(0, 1, n).astype(np.float32).astype(np.float64)
        ro = ob.step(a, u=u)
        gb.step(a, u=u, autoreset=False)
        po, co = ob.contacts(); pg, cg = gb.contacts(); pg, cg = pg.cpu().numpy(), cg.cpu().numpy()
        nt_g = gb.nticks.cpu().numpy()[:, 0]; obs_g = gb.obs.cpu().numpy().astype(np.float64)
        tg, sg = gb.terminated.cpu().numpy().astype(bool), gb.success.cpu().numpy().astype(bool)
        for i in range(n):
            tot += 1
            no, ng = int(ro['nticks'][i, 0]), int(nt_g[i])
            tdiff.append(abs(no - ng))
            if no != ng:
                l1 = ob.settle_trace(i)
                tick_margins.append(float(abs(l1[min(no, ng) - 1] - 5e-3)))
                continue
            same += 1
            obs_err.append(float(np.abs(obs_g[i] - ro['obs'][i]).max()))
            flags_ok += int(tg[i] == ro['terminated'][i] and sg[i] == ro['success'][i])
            so_ = sorted(tuple(p) for p in po[i, :co[i]].tolist()); sg_ = sorted(tuple(p) for p in pg[i, :cg[i]].tolist())
            if so_ == sg_:
                con_ok += 1
            else:
                con_margins.append(ob.contact_margin(i))
        done = ro['terminated'] | ro['truncated']
        if done.any():
            ob.reset(mask=done)
    gb.close()
    q = lambda x, p: float(np.quantile(x, p)) if len(x) else None
    return dict(env=name, precision=precision, cases=tot, same_ticks=same / tot, tick_diff_p90=q(tdiff, 0.9), tick_diff_max=int(max(tdiff)),
                tick_mismatch_margins=sorted(tick_margins)[-8:], tick_mismatch_margin_max=max(tick_margins) if tick_margins else None,
                flags_ok=flags_ok / max(same, 1), contacts_ok=con_ok / max(same, 1), contact_mismatch_margin_max=max(con_margins) if con_margins else None,
                contact_mismatches=len(con_margins), obs_err_p50=q(obs_err, 0.5), obs_err_p95=q(obs_err, 0.95), obs_err_max=q(obs_err, 1.0))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--envs', type=int, default=32)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--out', default='')
    ap.add_argument('--precisions', default='f64,f32')
    ap.add_argument('names', nargs='*')
    a = ap.parse_args()
    for name in (a.names or ENVS):
        for prec in a.precisions.split(','):
            m = run(name, prec, a.envs, a.steps)
            print(json.dumps(m), flush=True)
            if a.out:
                with open(a.out, 'a') as f:
                    f.write(json.dumps(m) + '\n')


if __name__ == '__main__':
    main()
