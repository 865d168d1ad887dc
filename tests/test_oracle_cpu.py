"""Known-answer and property tests of the CPU oracle's kinematics / dynamics / collision / solver (no GPU).

The rigid-body arithmetic has no reference fixture (pybullet is absent: "parity unpinned"), so what can be checked
here is physics the oracle must satisfy whatever Bullet's constants are: independent numpy kinematics from the
compiled URDF table, gravity torques from the potential energy, symmetric positive definite M^-1, closed-form
narrowphase cases, free fall against the integrator it documents, resting contact, gear coupling, determinism.
"""
import json
import math
import os

import numpy as np
import pytest

import oracle_lib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODEL = json.load(open(os.path.join(ROOT, 'urgym_b200', 'model', 'ur5_robotiq_85.json')))['links']
REST = [-1.5690622952052096, -1.5446774605904932, 1.343946009733127, -1.3708613585093699, -1.5707970583733368, 0.0009377758247187636]
ENVS = ['BallBalance', 'TwoBallsBalance', 'CubesPush', 'CubesGrasp', 'Golf', 'CubesPushDiff', 'GolfJoints']
SCENE_OF = {'CubesPushDiff': 'CubesPush', 'GolfJoints': 'Golf'}   # variants live in their parents' worlds


@pytest.fixture(scope='module', autouse=True)
def _built(built):
    return built


def _axis_rot(axis, ang):
    a = np.asarray(axis, float); a = a / np.linalg.norm(a)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + math.sin(ang) * K + (1 - math.cos(ang)) * K @ K


def numpy_fk(q):
    """Independent forward kinematics from the compiled link table: world (R, p) of every link frame."""
    R, P, dof = [None] * 19, [None] * 19, 0
    for l in MODEL:
        i, par = l['index'], l['parent']
        Rp, pp = (np.eye(3), np.zeros(3)) if par < 0 else (R[par], P[par])
        P[i] = pp + Rp @ np.array(l['jxyz'])
        R[i] = Rp @ np.array(l['jrot']).reshape(3, 3)
        if l['jtype'] == 1:
            R[i] = R[i] @ _axis_rot(l['axis'], q[dof]); dof += 1
    return R, P


def _quat_to_mat(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def _rand_q(rng):
    q = np.zeros(12)
    q[:6] = rng.uniform(-2.5, 2.5, 6)
    g = rng.uniform(0, 0.8)
    q[6:] = [g, -g, g, g, -g, g]
    return q


def test_fk_matches_independent_numpy_chain():
    rng = np.random.default_rng(1)
    for _ in range(20):
        q = _rand_q(rng)
        R, P = numpy_fk(q)
        out = O.fk(q)
        for i in range(19):
            np.testing.assert_allclose(out[i, :3], P[i], atol=1e-12)
            np.testing.assert_allclose(_quat_to_mat(out[i, 3:]), R[i], atol=1e-12)


def test_fk_rest_pose_landmarks():
    q = np.zeros(12); q[:6] = REST
    out = O.fk(q)
    np.testing.assert_allclose(out[7, :3], [0.110, -0.490, 0.510], atol=2e-3)   # ee_link at the rest pose
    assert abs(out[12, 2] - out[17, 2]) < 1e-3                                   # finger pads level
    assert np.linalg.norm(out[12, :3] - out[17, :3]) == pytest.approx(0.085 + 2 * 0.003175, abs=0.02)  # open gripper ~85 mm


def test_ik_reaches_reachable_targets():
    rng = np.random.default_rng(2)
    q = np.zeros(12); q[:6] = REST
    for _ in range(10):
        goal_q = q.copy(); goal_q[:6] += rng.uniform(-0.15, 0.15, 6)
        goal = O.fk(goal_q)[7]
        sol = q.copy()
        for _ in range(6):  # the reference re-issues IK every step; a few calls converge
            sol[:6] = O.ik(sol, goal)
        got = O.fk(sol)[7]
        assert np.linalg.norm(got[:3] - goal[:3]) < 2e-3
        assert min(np.linalg.norm(got[3:] - goal[3:]), np.linalg.norm(got[3:] + goal[3:])) < 5e-3


def test_ik_single_call_is_damped_and_bounded():
    q = np.zeros(12); q[:6] = REST
    far = O.fk(q)[7].copy(); far[:3] += [0.3, -0.2, 0.2]
    sol = O.ik(q, far)
    assert np.all(np.isfinite(sol)) and np.max(np.abs(sol - q[:6])) < 20 * math.pi / 4 + 1e-9
    same = O.ik(q, O.fk(q)[7])
    np.testing.assert_allclose(same, q[:6], atol=1e-9)   # already there: no motion


def test_minv_is_symmetric_positive_definite():
    rng = np.random.default_rng(3)
    for _ in range(5):
        Mi = O.minv(_rand_q(rng))
        np.testing.assert_allclose(Mi, Mi.T, atol=1e-9 * np.abs(Mi).max())
        assert np.linalg.eigvalsh(0.5 * (Mi + Mi.T)).min() > 0


def test_gravity_torques_equal_potential_gradient():
    """qdd(q, qd=0) = -M^-1 dV/dq, with V from the independent numpy kinematics."""
    rng = np.random.default_rng(4)

    def V(q, g=9.8):
        R, P = numpy_fk(q)
        return sum(l['mass'] * g * (P[l['index']] + R[l['index']] @ np.array(l['com']))[2] for l in MODEL)
    for _ in range(3):
        q = _rand_q(rng)
        grad = np.array([(V(q + h) - V(q - h)) / 2e-6 for h in np.eye(12) * 1e-6])
        qdd = O.forward_dynamics(q, np.zeros(12), gz=-9.8)
        np.testing.assert_allclose(qdd, -O.minv(q) @ grad, rtol=2e-4, atol=2e-3)


def test_forward_dynamics_no_gravity_no_motion():
    q = _rand_q(np.random.default_rng(5))
    np.testing.assert_allclose(O.forward_dynamics(q, np.zeros(12), gz=0.0), 0, atol=1e-9)


def test_euler_quaternion_round_trip_and_gimbal():
    rng = np.random.default_rng(6)
    for _ in range(50):
        e = rng.uniform([-3.1, -1.5, -3.1], [3.1, 1.5, 3.1])
        np.testing.assert_allclose(O.euler_from_quat(O.quat_from_euler(e)), e, atol=1e-9)
    e = O.euler_from_quat(O.quat_from_euler([0.3, math.pi / 2, 0.1]))
    assert e[1] == pytest.approx(math.pi / 2, abs=1e-3) and np.all(np.isfinite(e))


def _box(c, h, R=None):
    return np.concatenate([np.asarray(c, float), (np.eye(3) if R is None else R).reshape(-1), np.asarray(h, float)])


def test_box_box_face_contact_gives_four_points():
    pts = O.box_box(_box([0, 0, 0.0199], [0.02] * 3), _box([0, 0, -0.25], [0.5, 1.0, 0.25]), 0.02)
    assert len(pts) == 4
    np.testing.assert_allclose(pts[:, 6:9], [[0, 0, 1]] * 4, atol=1e-12)     # normal on B pointing to A
    np.testing.assert_allclose(pts[:, 9], -0.0001, atol=1e-9)               # 0.1 mm penetration
    assert sorted(map(tuple, np.round(np.abs(pts[:, 3:5]), 6).tolist())) == [(0.02, 0.02)] * 4


def test_box_box_separated_and_rotated():
    assert len(O.box_box(_box([0, 0, 1.0], [0.02] * 3), _box([0, 0, -0.25], [0.5, 1.0, 0.25]), 0.02)) == 0
    R = _axis_rot([1, 0, 0], math.pi / 4)                                      # edge down
    pts = O.box_box(_box([0, 0, 0.02 * math.sqrt(2) - 0.001], [0.02] * 3, R), _box([0, 0, -0.25], [0.5, 1.0, 0.25]), 0.02)
    assert 1 <= len(pts) <= 4
    assert pts[:, 9].min() == pytest.approx(-0.001, abs=1e-6)
    np.testing.assert_allclose(pts[:, 6:9], [[0, 0, 1]] * len(pts), atol=1e-9)


def test_urand_is_24_bit_uniform_and_keyed():
    L = O.lib()
    v = np.array([L.urgo_urand(7, e, 1, d) for e in range(200) for d in range(4)])
    assert v.min() >= 0 and v.max() < 1 and np.all(v * 2 ** 24 == np.round(v * 2 ** 24))
    assert abs(v.mean() - 0.5) < 0.05 and len(set(v.tolist())) > 790
    assert L.urgo_urand(7, 3, 1, 0) != L.urgo_urand(8, 3, 1, 0) != L.urgo_urand(7, 3, 2, 0)


RESET_TICKS = {'BallBalance': (170, 180), 'TwoBallsBalance': (170, 180), 'CubesPush': (130, 155), 'CubesGrasp': (11, 11), 'Golf': (180, 192)}


@pytest.mark.parametrize('name', ENVS)
def test_reset_is_deterministic_and_layout_independent(name):
    a, b = O.OracleBatch(name, 4, seed=9), O.OracleBatch(name, 2, seed=9, offset=2)
    oa, ta = a.reset(); ob, tb = b.reset()
    np.testing.assert_array_equal(oa[2:], ob)          # env i of a shard at offset k == env i+k of the whole batch
    np.testing.assert_array_equal(ta[2:], tb)
    lo, hi = RESET_TICKS[SCENE_OF.get(name, name)]
    assert np.all((ta >= lo) & (ta <= hi)), ta          # 10 arm ticks + the settle loops (regression band)
    if name != 'CubesGrasp':
        assert len({tuple(r) for r in np.round(oa, 9).tolist()}) == 4   # different draws per env
    a2 = O.OracleBatch(name, 4, seed=9)
    np.testing.assert_array_equal(a2.reset()[0], oa)


@pytest.mark.parametrize('name', ENVS)
def test_reset_leaves_a_sane_world(name):
    ob = O.OracleBatch(name, 2, seed=4)
    full_name, name = name, SCENE_OF.get(name, name)
    obs, _ = ob.reset()
    st = ob.get_state()
    facts = O.model_facts()
    assert np.all(np.isfinite(st)) and np.all(np.isfinite(obs))
    for s in st:
        q = s[0:12]
        tol = 2e-2 if name in ('CubesPush', 'CubesGrasp') else 0.25              # fingers clamped on an object strain the gears
        for child, ratio in zip([7, 8, 9, 10, 11], facts['gear_ratio']):        # mimic gears hold
            assert abs(q[6] + ratio * q[child]) < tol
        np.testing.assert_allclose(s[0:6], np.clip(s[24:30], -math.pi, math.pi), atol=0.08)  # arm near its (limit-clipped) targets
        flags = int(s[31])
        for b in range(3):
            if flags & (16 << b):
                assert np.linalg.norm(s[41 + 13 * b: 45 + 13 * b]) == pytest.approx(1.0, abs=1e-9)
    if name in ('CubesPush', 'CubesGrasp'):
        if name == 'CubesGrasp':   # reset settles on the (idle) arm after one tick; the cube is still falling
            ob.tick(240); st = ob.get_state()
        zs = [s[40 + 13 * b] for s in st for b in range(2 if name == 'CubesPush' else 1)]
        np.testing.assert_allclose(zs, 0.02, atol=2.5e-3)                        # 4 cm cubes at rest on the surface
    if name in ('BallBalance', 'TwoBallsBalance'):
        assert np.all(obs[:, 5] < 0)                                             # ball above the paddle
    assert obs.shape[1] == ob.od and full_name in O.ENV_TYPES


def test_free_fall_follows_the_documented_integrator():
    """CubesGrasp: gravity -10, v += dt (g - v (k + k|v|)), x += dt v (semi-implicit Euler, damping k = 0.04)."""
    ob = O.OracleBatch('CubesGrasp', 1)
    ob.reset()
    s = ob.get_state()[0]
    s[38:41] = [0.6, 0.3, 1.0]; s[41:45] = [0, 0, 0, 1]; s[45:51] = 0
    ob.set_state(s)
    z, v, dt = 1.0, 0.0, 1 / 240
    for _ in range(60):
        v += dt * (-10.0 - v * (0.04 + 0.04 * abs(v))); z += dt * v
    ob.tick(60)
    got = ob.get_state()[0]
    assert got[40] == pytest.approx(z, abs=1e-9) and got[47] == pytest.approx(v, abs=1e-9)
    np.testing.assert_allclose(got[38:40], [0.6, 0.3], atol=1e-12)


def test_zero_action_keeps_the_ball_on_the_paddle():
    ob = O.OracleBatch('BallBalance', 2, seed=1)
    ob.reset()
    for _ in range(5):
        r = ob.step(np.zeros((2, 3)))
        assert not r['terminated'].any() and np.all(r['reward'] == 1.0)
        assert np.all(r['nticks'][:, 0] >= 1) and np.all(r['nticks'][:, 0] <= 480)


def test_autoreset_returns_first_observation_of_next_episode():
    ob = O.OracleBatch('CubesGrasp', 2, seed=3)
    ob.reset()
    a = np.tile(np.array([1, 0] + REST + [0], float), (2, 1))
    last = None
    for k in range(20):
        last = ob.step(a, u=np.array([0.2, 0.2]), autoreset=True)
    assert last['truncated'].all() and not last['terminated'].any()      # 20 = max_episode_steps
    st = ob.get_state()
    assert np.all(st[:, 32] == 0) and np.all(st[:, 33] == 2)             # counters restarted, second episode
    assert np.all(last['nticks'][:, 1] >= 11)                            # reset ticks reported separately
    assert not np.allclose(last['final_obs'], last['obs'])
    nxt = ob.step(a, u=np.array([0.2, 0.2]), autoreset=True)
    assert not nxt['truncated'].any()


def test_scripted_grasp_reaches_every_sub_goal_through_the_physics():
    """approach -> close -> raise on CubesGrasp (tests/scripted_grasp.py): +5 approached, +5 grasped, +10 raised with is_success,
    none of which a random policy ever sees (env_cubes_grasp_v0.py:150-168, 206-233)."""
    import scripted_grasp as G
    n = 4
    ob = O.OracleBatch('CubesGrasp', n, seed=11)
    ob.reset()
    log = G.run(ob, n)
    bits = np.array([r['obs'][:, :2] for _, r in log])            # [step, env, (approached, grasped)]
    rew = np.array([r['reward'] for _, r in log])
    assert (bits[0] == 0).all() and (bits[-1] == 1).all()
    first_app = (bits[:, :, 0] == 1).argmax(0); first_grasp = (bits[:, :, 1] == 1).argmax(0)
    assert (first_app == 2).all() and (first_grasp == 3).all()     # down-around-the-cube step, then the closing step
    assert np.allclose(rew[2], 5.15, atol=0.1) and np.allclose(rew[3], 4.9, atol=1e-9)
    last = log[-1][1]
    assert last['terminated'].all() and last['success'].all() and np.allclose(last['reward'], 9.9) and (last['obs'][:, 12] > 0.2).all()
    assert len(log) <= 20                                          # inside the env's 20-step TimeLimit


def test_recorded_episode_statistics_are_of_this_oracle():
    """tests/golden/episode_stats_oracle.json (the oracle's side of the GPU suite's 1e4-episode test, recorded by
    tests/golden/make_episode_stats.py) must be regenerated when oracle/ changes: a small fresh sample of the same policy has to
    land on the recorded means (CubesPush: every episode runs its 50 steps; return -5.19 with a per-episode spread of 0.09)."""
    want = json.load(open(os.path.join(ROOT, 'tests', 'golden', 'episode_stats_oracle.json')))['CubesPush']
    n = 48
    ob = O.OracleBatch('CubesPush', n, seed=41)
    ob.reset()
    rng = np.random.default_rng(11)
    ret, eps = np.zeros(n), []
    for k in range(50):
        a = rng.uniform(-0.1, 0.1, (n, 3)).astype(np.float32).astype(np.float64); rng.uniform(0, 1, n)
        r = ob.step_parallel(a, autoreset=True)
        ret += r['reward']
        done = r['terminated'] | r['truncated']
        eps += ret[done].tolist(); ret[done] = 0
    assert len(eps) == n
    se = np.sqrt(want['var']['return'] / n)
    assert abs(np.mean(eps) - want['mean']['return']) < 5 * se + 0.01, (np.mean(eps), want['mean'])
    assert want['mean']['length'] == 50.0 and want['episodes'] >= 10000
