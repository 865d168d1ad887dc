"""Rollout trace format (urgym_b200/trace.py): record / save / load / replay, on CPU with the oracle on both sides."""
import numpy as np
import pytest

import oracle_lib as O
from urgym_b200 import trace as T


def test_trace_round_trip_and_exact_replay(built, tmp_path):
    rng = np.random.default_rng(0)
    a = O.OracleBatch('CubesPush', 3, seed=4); a.reset()
    actions = rng.uniform(-0.1, 0.1, (4, 3, 3)).astype(np.float32)
    tr = T.record(a, 'CubesPush', actions)
    assert tr['obs'].shape == (4, 3, 9) and tr['state'].shape == (4, 3, 96) and tr['pairs'].shape == (4, 3, 64, 2)
    path = str(tmp_path / 'push.npz')
    T.save(path, tr)
    back = T.load(path)
    assert back['env'] == 'CubesPush' and all(np.array_equal(back[k], tr[k]) for k in T.FIELDS)
    b = O.OracleBatch('CubesPush', 3, seed=99); b.reset()          # another instance: everything comes from the trace
    for resync in (True, False):
        d = T.replay(b, back, resync=resync)
        assert d['obs_err'].max() == 0 and d['state_err'].max() == 0 and d['reward_err'].max() == 0
        assert d['same_ticks'].min() == 1 and d['same_flags'].min() == 1 and d['same_contacts'].min() == 1


def test_trace_rejects_other_versions(built, tmp_path):
    a = O.OracleBatch('BallBalance', 1); a.reset()
    tr = T.record(a, 'BallBalance', np.zeros((1, 1, 3), np.float32))
    tr['format_version'] = 7
    T.save(str(tmp_path / 'x.npz'), tr)
    with pytest.raises(ValueError):
        T.load(str(tmp_path / 'x.npz'))
