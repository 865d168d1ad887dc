"""The reference's shipped SB3 actors, trained on pybullet, replayed on the CPU oracle (tools/replay_actors.py).

The only end-to-end signal about the rigid-body arithmetic that exists without pybullet: a policy trained on Bullet's
dynamics must still solve the task on dynamics that behave like Bullet's. CubesPush does (about 9 in 10 episodes
against 10 in 10 on pybullet); Golf and GolfJoints do much less often — recorded in profiles/r01_actor_replay.json and
DESIGN.md §3 as the measured distance of the restatement from Bullet, not asserted here.
Needs /root/reference (the checkpoints are not copied into the repo); skipped where it is absent."""
import os
import sys

import pytest

REF = os.environ.get('URGYM_REFERENCE', '/root/reference')
CKPT = os.path.join(REF, 'policies_final', 'cubes_push_v0', 'SAC_policy.zip')


@pytest.mark.skipif(not os.path.exists(CKPT), reason='reference checkpoints not present on this machine')
def test_pybullet_trained_cubes_push_actor_solves_the_task_on_the_oracle(built):
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tools'))
    import replay_actors as R
    act, hist = R.load_actor(CKPT)
    r = R.rollout('CubesPush', act, episodes=64, seed=7, threads=min(8, os.cpu_count() or 1))
    assert r['success_rate'] >= 0.55, r          # measured 0.80 over 200 episodes; pybullet: 100 / 100 in the checkpoint's Monitor buffer
    assert r['mean_length'] <= 30, r             # pybullet: 8.1 steps; a random policy never succeeds within the 50-step limit
