"""Box-v0 (urgym/envs/env_box_v0.py): the oracle's env layer against fixtures recorded from the reference's own
BoxManipulation class (tests/golden/reference_box_logic.json, made by tests/golden/make_box_golden.py with the scripted
pybullet stand-in), and the oracle's model of the box (urgym/urdf/skew-box-button.urdf) against facts checkable without Bullet.
The box's rigid-body arithmetic is PARITY UNPINNED like the rest; its four cylinders collide as square prisms."""
import json
import os

import numpy as np
import pytest

import oracle_lib as O

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), 'golden', 'reference_box_logic.json')))


@pytest.fixture(scope='module', autouse=True)
def _built(built):
    return built


def test_registry_spaces_and_construction():
    from urgym_b200 import envs as E
    assert GOLD['registry']['max_episode_steps'] == E.SPECS['URGym/Box-v0'][1] == O.OracleBatch('Box', 1).max_steps == 200
    assert GOLD['registry']['kwargs'] == {'button_touch_mode': 'any'}
    o, a = E._spaces('URGym/Box-v0')
    g = GOLD['spaces']
    np.testing.assert_array_equal(o.low, np.array(g['obs_low'])); np.testing.assert_array_equal(o.high, np.array(g['obs_high']))
    np.testing.assert_array_equal(a.low, np.array(g['act_low'], np.float32)); np.testing.assert_array_equal(a.high, np.array(g['act_high'], np.float32))
    assert str(o.dtype) == g['obs_dtype'] and str(a.dtype) == g['act_dtype']
    # gravity -10, plane at the origin, robot base at (0, 0.5, 0), box with a fixed base at the origin, merged fixed links + self-collision
    c = GOLD['construction']
    assert c[0] == ['setGravity', [0, 0, -10], {}] and c[1][1][0] == 'plane.urdf' and c[2][1][1] == [0, 0.5, 0] == GOLD['robot_base']
    assert c[3][1][0] == 'skew-box-button.urdf' and c[3][1][1] == [0.0, 0.0, 0.0] and c[3][2]['useFixedBase'] and c[3][2]['flags'] == 512 | 8


def test_reset_command_sequence():
    """robot.reset() (rest pose, 10 ticks, gripper open), move_ee to the CURRENT pose, reset_box(): joints to 0, button position
    motor with force 1, door motor off (env_box_v0.py:172-185)."""
    r = GOLD['reset']
    ops = [c['op'] for c in r['commands']]
    assert ops[:4] == ['reset_arm', 'ticks', 'gripper', 'ik'] and r['commands'][1]['n'] == 10
    assert r['box_calls'] == [['resetJointState', [0, 0], {}], ['resetJointState', [1, 0], {}], ['setJointMotorControl2', [0, 2], {'force': 1}],
                              ['setJointMotorControl2', [1, 0], {'force': 0}]]
    ob = O.OracleBatch('Box', 1, seed=3)
    ob.set_trace(True)
    obs, nt = ob.reset()
    assert nt[0] == 10
    tr = ob.trace()
    assert [int(e[0]) for e in tr[:4]] == [7, 4, 2, 1]                      # EV_RESET_ARM, EV_TICKS(10), EV_GRIPPER, EV_IK
    np.testing.assert_allclose(tr[3][1:8], r['commands'][3]['pos'] + r['commands'][3]['orn'], atol=1e-9)   # IK target = current pose
    np.testing.assert_allclose(obs[0], r['obs'], atol=1e-9)
    s = ob.get_state()[0]
    np.testing.assert_allclose(s[38:45], [0.075, 0, 0.1, 0, 0, 0, 1]); np.testing.assert_allclose(s[51:58], [0, 0, 0.07, 0, 0, 0, 1])


@pytest.mark.parametrize('i', range(len(GOLD['step'])))
def test_step_logic_matches_the_reference(i):
    c = GOLD['step'][i]
    ob = O.OracleBatch('Box', 1, seed=1)
    ob.reset()
    ob.set_state(np.array(c['state'])[None])
    ob.set_trace(True)
    pro = ob.prologue(np.array(c['action'], dtype=np.float64))
    tr = ob.trace()
    ik, grip = [x for x in c['commands'] if x['op'] == 'ik'][0], [x for x in c['commands'] if x['op'] == 'gripper'][0]
    ev_ik = [e for e in tr if int(e[0]) == 1][0]; ev_gr = [e for e in tr if int(e[0]) == 2][0]
    np.testing.assert_allclose(ev_ik[1:8], ik['pos'] + ik['orn'], atol=1e-7)   # pose + action[:7]; the action is float32 in the reference
    assert ev_gr[1] == pytest.approx(grip['target'], abs=1e-7)
    assert [x for x in c['commands'] if x['op'] == 'ticks'][0]['n'] == 120
    ro = ob.epilogue(np.zeros((0, 2), np.int32), reward_so_far=pro['reward'])
    on_threshold = 'exactly' in c['label']
    if not on_threshold:
        assert ro['reward'] == pytest.approx(c['reward'], abs=1e-12)
        assert ro['terminated'] == c['terminated'] and ro['success'] == c['info']['is_success']
        fl = int(ob.get_state()[0, 31])
        assert bool(fl & 4) == c['flags_after']['box_opened'] and bool(fl & 8) == c['flags_after']['button_pressed']
        assert ro['terminated'] == c['flags_after']['box_closed']
    np.testing.assert_allclose(ro['obs'], c['obs'], atol=1e-9)


def test_box_model_facts():
    """Checkable without Bullet: the door rests closed on its joint limit and the button at its upper stop with nothing touching
    them; a door opened past the vertical falls open onto its upper limit (2 rad), one below it falls shut; a pressed button
    comes back up under its 1 N position motor."""
    ob = O.OracleBatch('Box', 1, seed=5)
    ob.reset()
    idle = np.zeros((1, 8)); idle[0, 7] = 0.085

    def run(door, button, steps=3):
        s = ob.get_state()
        s[0, 41:45] = [0, np.sin(door / 2), 0, np.cos(door / 2)]; s[0, 45:51] = 0; s[0, 53] = 0.07 + button; s[0, 58:64] = 0
        ob.set_state(s)
        for _ in range(steps):
            r = ob.step(idle)
            assert r['nticks'][0, 0] == 120
        s = ob.get_state()[0]
        return 2 * np.arctan2(s[42], s[44]), s[53] - 0.07
    d, b = run(0.0, 0.0)
    assert abs(d) < 2e-3 and abs(b) < 1e-4
    d, b = run(1.9, -0.04, steps=6)
    assert abs(d - 2.0) < 5e-3 and abs(b) < 1e-4          # past 90 degrees the door falls open; the button returns
    d, b = run(1.2, 0.0, steps=14)
    assert abs(d) < 5e-3                                   # below it the door falls shut (onto the button and its limit)


def test_scripted_policy_solves_the_task_on_the_oracle():
    """tests/scripted_box.py: grasp the handle bar, lift the door along its arc past 1.9 rad (+1), let go, press the button with
    the closed gripper (+1), push the door back over its tipping point so that it falls shut (+1, terminated, is_success) — all
    through the physics, inside the 200-step limit."""
    import scripted_box as SB
    ob = O.OracleBatch('Box', 1, seed=1)
    ob.reset()
    log = SB.run(lambda a: ob.step(a), 1)
    wins = [k for k, r in enumerate(log) if r['reward'][0] > 0.5]
    assert len(wins) == 3 and len(log) < 200, (wins, len(log))
    assert log[-1]['terminated'][0] and log[-1]['success'][0] and not any(r['terminated'][0] for r in log[:-1])
    assert sum(r['reward'][0] for r in log) == pytest.approx(3.0 - 0.1 * len(log), abs=1e-9)
