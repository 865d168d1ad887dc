#!/usr/bin/env python
"""BASELINE.json config 0 ("URGym/BallBalance-v0 single env, ... random actions 1000 steps (reference correctness trace)") as a
committed artefact: tests/golden/config0_ballbalance_trace.npz.

SURVEY.md section 8d "Config 1": one env, 1000 gym steps, actions ~ U(-0.1, 0.1)^3 float32 from torch.Generator(seed=0),
auto-reset on terminated or at 500 steps, reset draws (ball dx ~ U(-0.03, 0.03), dy ~ U(-0.08, 0.02),
env_ball_balance_v0.py:193-194) from random.Random(0), injected as uniforms. pybullet is not installable here, so the trace is
recorded on the fp64 CPU restatement (oracle/); on a machine with pybullet the same script's inputs (actions, draws) drive the
reference itself and urgym_b200.trace.replay compares. tests/test_trace.py replays it on the oracle (CPU) and
tests/test_gpu_parity.py on both CUDA builds.

    python tests/golden/make_config0_trace.py
"""
import os
import random
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import oracle_lib as O  # noqa: E402
from urgym_b200 import trace as T  # noqa: E402

STEPS = 1000


def inputs():
    g = torch.Generator(); g.manual_seed(0)
    actions = ((torch.rand(STEPS, 1, 3, generator=g) * 2 - 1) * 0.1).numpy().astype(np.float32)
    r = random.Random(0)
    seq = np.array([[r.random(), r.random(), 0.0, 0.0] for _ in range(STEPS + 1)])   # the k-th reset consumes the k-th (dx, dy) pair
    return actions, seq[:1], (lambda t, resets: seq[1 + resets])


def main():
    actions, first, draws = inputs()
    b = O.OracleBatch('BallBalance', 1, seed=0)
    b.reset(draws=first)
    tr = T.record(b, 'BallBalance', actions, draws=draws)
    tr['first_draws'] = first
    T.save(os.path.join(HERE, 'config0_ballbalance_trace.npz'), tr)
    ep = int((tr['terminated'] | tr['truncated']).sum())
    print('steps', STEPS, 'episodes finished', ep, 'mean settle ticks', tr['nticks'][:, 0, 0].mean(), 'reward sum', tr['reward'].sum())


if __name__ == '__main__':
    main()
