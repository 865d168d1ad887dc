"""The reference's shipped SB3 actors, trained on pybullet, replayed on the CPU oracle (tools/replay_actors.py).

The only end-to-end signal about the rigid-body arithmetic that exists without pybullet: a policy trained on Bullet's
dynamics must still solve the task on dynamics that behave like Bullet's. The actor weights travel as a fixture
(tests/golden/actors.npz, extracted by tests/golden/make_actors.py with torch.load(weights_only=True)); the targets are the
Monitor statistics of the actors' last 100 training episodes on pybullet (tests/golden/actors.json). Measured with 200
episodes each (profiles/r02_geometry_ablation.txt): Golf 95 % (pybullet 99 %), CubesPush 77 % (100 %), GolfJoints 71 % (98 %).
The bars here are regression bars on small samples; the GPU suite repeats the replay at 1e4 episodes on the device."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tools'))


@pytest.mark.parametrize('name,episodes,min_success,max_length', [('CubesPush', 64, 0.55, 30), ('Golf', 48, 0.80, 30)])
def test_pybullet_trained_actor_solves_the_task_on_the_oracle(built, name, episodes, min_success, max_length):
    import replay_actors as R
    r = R.rollout_oracle(name, episodes, seed=7, threads=min(8, os.cpu_count() or 1))
    assert r['success_rate'] >= min_success, r
    assert r['mean_length'] <= max_length, r


def test_actor_fixture_matches_its_description():
    import actors as A
    import numpy as np
    for env, m in A.META.items():
        w = A.weights(env)
        assert w['W0'].shape[1] == m['obs_dim'] and w['Wm'].shape[0] == m['act_dim']
        a = A.numpy_actor(env)(np.zeros((3, m['obs_dim'])))
        assert a.shape == (3, m['act_dim']) and np.all(np.abs(a) <= m['action_scale'] + 1e-12)
