"""Parity of the CUDA path (through the C ABI) with the CPU oracle and with the reference-recorded fixtures. B200 only.

Bars (fp32 device arithmetic against the fp64 oracle; tolerances stated at each assert):
  * kinematics (FK / IK / gripper centre): <= 1e-5 absolute.
  * reset(): tick counts EXACT, first observation <= 1e-4.
  * one physics tick from an identical state: positions <= 2e-5, joint / linear velocities <= 2e-3, free-body spin
    <= 2e-2 rad/s (one fp32 PGS solve; a 3 mm ball turns a 6e-5 m/s difference into 2e-2 rad/s).
  * one gym step from an identical state (re-synchronised every step so that chaotic growth does not accumulate), integer
    outputs (settle-loop trip count, terminated / success, set of contact pairs):
      - parity build (fp64 kernels, liburgym_b200_f64.so): EQUAL to the oracle on every (env, step) pair for the ball and
        cube-pushing scenes; for CubesGrasp and the Golf scenes, whose one-step map amplifies a 1e-13 perturbation of the
        start state into a different trip count on 0.1 % / 6 % / 10 % of the steps ON THE ORACLE ITSELF (measured in the
        test), at least as often as the oracle agrees with its own perturbed copy (minus 3 points of sampling slack).
        This is what shows that kernel and oracle run the same algorithm.
      - product build (fp32): every trip-count mismatch is recorded with the oracle's decision margin (|summed joint motion
        - 5e-3| at the tick where one side stopped); agreement must be at least the oracle's own agreement under a perturbation
        of the start state of the size fp32 adds per tick (1e-6) minus 8 points, and >= 95 % on the ball / cube-pushing scenes; where the trip
        counts agree, observations agree to 5e-3 (p95), flags on >= 99 % and contact-pair sets on >= 95 % (all eight envs).
  * env logic (commands, predicates, reward, termination, observation) against tests/golden/reference_env_logic.json,
    i.e. against the reference's own step(): obs <= 2e-5, reward <= 2e-4, flags exact (fixture cases that sit exactly
    on a threshold are excluded — fp32 may land on either side).
  * Box-v0: the same reset / tick / step checks; env logic against tests/golden/reference_box_logic.json (the reference's own
    BoxManipulation.step()); a button press driven through the physics with equal sub-goal bits and rewards at every step.
  * >= 1e4 episodes per side under one random policy: mean return / length / success rate within 4 standard errors.
  * full-size runs (4096 / 16384 envs): finite, unit quaternions, episode accounting, time-limit truncation,
    layout independence (bit-exact), host-buffer entry == device entry (bit-exact).
"""
import json
import os

import numpy as np
import pytest
import torch

import oracle_lib as O

pytestmark = pytest.mark.gpu
ENVS = ['BallBalance', 'TwoBallsBalance', 'CubesPush', 'CubesGrasp', 'Golf', 'CubesPushDiff', 'GolfJoints', 'Box']
SCENE_OF = {'CubesPushDiff': 'CubesPush', 'GolfJoints': 'Golf'}   # variants live in their parents' worlds
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'gpurun_out')


@pytest.fixture(scope='module', autouse=True)
def _need_gpu(built):
    if not torch.cuda.is_available():
        pytest.fail('no CUDA device: the gpu tests must run on the B200 box')
    return True


def _batch(*a, **k):
    from urgym_b200.batch import EnvBatch
    return EnvBatch(*a, **k)


def _np(t):
    return t.detach().cpu().numpy()


def _actions(rng, name, n):
    if name == 'GolfJoints':
        return rng.uniform(-0.1, 0.1, (n, 6)).astype(np.float32).astype(np.float64)
    if name == 'CubesGrasp':
        # the env's own action box (env_cubes_grasp_v0.py:74), which is also what bench.py samples: [0,1]^2 x [-pi,pi]^6 x [0,1]
        return np.concatenate([rng.uniform(0, 1, (n, 2)), rng.uniform(-np.pi, np.pi, (n, 6)), rng.uniform(0, 1, (n, 1))], 1).astype(np.float32).astype(np.float64)
    if name == 'Box':
        # the env's own action box (env_box_v0.py:44): [-0.05, 0.05]^3 x [-pi/10, pi/10]^4 x [0, 0.085], biased down towards the box
        a = np.concatenate([rng.uniform(-0.05, 0.05, (n, 3)), rng.uniform(-np.pi / 10, np.pi / 10, (n, 4)) * 0.2, rng.uniform(0, 0.085, (n, 1))], 1)
        a[:, 2] = -np.abs(a[:, 2])
        return a.astype(np.float32).astype(np.float64)
    return rng.uniform(-0.1, 0.1, (n, 3)).astype(np.float32).astype(np.float64)


def _record(name, payload):
    try:
        os.makedirs(OUT, exist_ok=True)
        with open(os.path.join(OUT, 'parity_metrics.jsonl'), 'a') as f:
            f.write(json.dumps(dict(test=name, **payload)) + '\n')
    except OSError:
        pass


# ------------------------------------------------------------------------------------------------ kinematics
def test_kinematics_match_oracle():
    from urgym_b200.batch import forward_kinematics, gripper_center, inverse_kinematics
    rng = np.random.default_rng(0)
    q = rng.uniform(-1.5, 1.5, (256, 12)); q[:, 6:] = rng.uniform(0, 0.8, (256, 1)) * np.array([1, -1, 1, 1, -1, 1])
    q = q.astype(np.float32).astype(np.float64)
    fk_g = _np(forward_kinematics(q)).astype(np.float64)
    fk_o = np.stack([O.fk(x) for x in q])
    assert np.abs(fk_g[..., :3] - fk_o[..., :3]).max() < 1e-5
    assert np.abs(fk_g[..., 3:] - fk_o[..., 3:]).max() < 1e-5                      # same sign convention too
    tg = fk_o[:, 7].copy(); tg[:, :3] += rng.uniform(-0.1, 0.1, (256, 3)); tg = tg.astype(np.float32).astype(np.float64)
    ik_g = _np(inverse_kinematics(q, tg)); ik_o = np.stack([O.ik(q[i], tg[i]) for i in range(256)])
    assert np.abs(ik_g - ik_o).max() < 1e-5
    gc_g = _np(gripper_center(q)); gc_o = np.stack([O.gripper_center(x) for x in q])
    assert np.abs(gc_g - gc_o).max() < 1e-5


# ------------------------------------------------------------------------------------------- reset / tick / step
@pytest.mark.parametrize('precision', ['f64', 'f32'])
@pytest.mark.parametrize('name', ENVS)
def test_reset_ticks_exact_and_first_observation(name, precision):
    n = 32
    ob, gb = O.OracleBatch(name, n, seed=3), _batch(name, n, seed=3, precision=precision)
    oo, ont = ob.reset()
    go = _np(gb.reset()).astype(np.float64)
    gt = _np(gb.reset_ticks)
    err = np.abs(go - oo).max()
    _record('reset', dict(env=name, precision=precision, obs_err=float(err), ticks_oracle=int(ont[0]), ticks_gpu=int(gt[0])))
    if precision == 'f64' or SCENE_OF.get(name, name) != 'Golf':
        np.testing.assert_array_equal(gt, ont)
    else:
        # fp32 on the Golf scenes: the second settle loop ends while the pinched 3 kg club still creeps in the fingers, so the
        # summed joint motion hovers at the 5e-3 threshold for several ticks (the fp64 build above is exact)
        assert np.abs(gt.astype(int) - ont.astype(int)).max() <= 6, (gt, ont)
    # fp64 build: round-off, except CubesGrasp, whose observation exposes the gripper opening: its reset ends one tick after the
    # open command, with the gear-coupled finger joints (inertias 1e-6..1e-8 kg m^2 behind a 4 kg arm) moving at full speed, and
    # the kernel's factorised mass matrix and the checker's articulated-body recursion round that ill-conditioned block
    # differently (measured: finger joints 1e-7 rad, opening 2e-7; every arm coordinate <= 2e-8)
    # (Box: the reset's IK runs to the CURRENT pose, as CubesGrasp's does: the rotation error goes through acos near 1)
    f64_bar = 1e-6 if name in ('CubesGrasp', 'Box') else 1e-9
    assert err < (f64_bar if precision == 'f64' else (2e-3 if SCENE_OF.get(name, name) == 'Golf' else 1e-4)), err
    # seeded draws come from the same counter-based generator on both sides: the randomised spawn coordinates agree
    assert np.isfinite(go).all()


@pytest.mark.parametrize('name', ENVS)
def test_single_tick_from_identical_state(name):
    n = 16
    ob, gb = O.OracleBatch(name, n, seed=5), _batch(name, n, seed=5)
    ob.reset()
    rng = np.random.default_rng(1)
    ob.step(_actions(rng, name, n), u=rng.uniform(0, 1, n))                          # a state with contacts and motion
    s0 = ob.get_state().astype(np.float32)
    gb.set_state(torch.tensor(s0)); ob.set_state(s0.astype(np.float64))
    ob.tick(1); gb.tick(1)
    so, sg = ob.get_state(), _np(gb.get_state()).astype(np.float64)
    pos = list(range(0, 12)) + [38 + 13 * b + k for b in range(3) for k in range(7)]
    vel = list(range(12, 24)) + [38 + 13 * b + k for b in range(3) for k in range(7, 10)]
    ang = [38 + 13 * b + k for b in range(3) for k in range(10, 13)]   # spin of the free bodies: v / r with r down to 3 mm
    ep, ev, ea = np.abs(sg[:, pos] - so[:, pos]).max(), np.abs(sg[:, vel] - so[:, vel]).max(), np.abs(sg[:, ang] - so[:, ang]).max()
    _record('single_tick', dict(env=name, pos_err=float(ep), vel_err=float(ev), spin_err=float(ea)))
    # The Golf scenes hold a 3 kg club in a mu = 10 pinch: the finger joints (inertias down to 1e-7 kg m^2) take the fp32
    # solve's error as velocity (measured 1.6e-2 rad/s on two finger joints, 6.6e-5 rad after the tick; arm joints 3e-4 / 1e-6).
    # The fp64 build of the same kernel leaves 6e-12 rad/s there (tools/diag_tick.py), so this is the scalar width.
    golf = SCENE_OF.get(name, name) == 'Golf'
    assert ep < (2e-4 if golf else 2e-5) and ev < (5e-2 if golf else 2e-3) and ea < 2e-2


def _pairs_of(points, count):
    return {tuple(p) for p in points[:count].tolist()}


def _resync_run(name, precision, n=32, steps=None, perturb=0.0):
    """50 (CubesGrasp, Box: 25) re-synchronised gym steps of `n` envs on the oracle and on the CUDA build `precision` (or, precision=None, on a
    second oracle whose start state is perturbed by `perturb`: the oracle's own sensitivity). Returns the agreement metrics."""
    steps = steps or (25 if name in ('CubesGrasp', 'Box') else 50)   # 120-180 ticks per step on the CPU oracle: half the steps, 800 cases
    ob = O.OracleBatch(name, n, seed=7)
    gb = _batch(name, n, seed=7, precision=precision) if precision else O.OracleBatch(name, n, seed=7)
    ob.reset(); gb.reset()
    rng = np.random.default_rng(2); prng = np.random.default_rng(99)
    npdt = np.float32 if precision == 'f32' else np.float64
    same_ticks = total = flags_ok = contacts_ok = 0
    obs_err, rew_err, t_o, t_g, tick_margin = [], [], [], [], []
    for k in range(steps):
        s = ob.get_state().astype(npdt)
        ob.set_state(s.astype(np.float64))
        if precision:
            gb.set_state(torch.tensor(s))
        else:
            sp = s.astype(np.float64).copy(); sp[:, :24] += perturb * prng.standard_normal((n, 24)); gb.set_state(sp)
        a, u = _actions(rng, name, n), rng.uniform(0, 1, n).astype(np.float32).astype(np.float64)
        ro = ob.step_parallel(a, u=u)
        po, co = ob.contacts()
        if precision:
            gb.step(a, u=u, autoreset=False)
            sets_g = gb.contact_pairs()
            nt_g, obs_g, rew_g = _np(gb.nticks)[:, 0], _np(gb.obs).astype(np.float64), _np(gb.reward).astype(np.float64)
            tg, sg = _np(gb.terminated).astype(bool), _np(gb.success).astype(bool)
        else:
            rg = gb.step_parallel(a, u=u)
            pg, cg = gb.contacts(); sets_g = [_pairs_of(pg[i], cg[i]) for i in range(n)]
            nt_g, obs_g, rew_g, tg, sg = rg['nticks'][:, 0], rg['obs'], rg['reward'], rg['terminated'], rg['success']
        for i in range(n):
            total += 1
            no, ng = int(ro['nticks'][i, 0]), int(nt_g[i])
            t_o.append(no); t_g.append(ng)
            if ng != no:
                tick_margin.append(float(abs(ob.settle_trace(i)[min(no, ng) - 1] - 5e-3)))
                continue
            same_ticks += 1
            obs_err.append(np.abs(obs_g[i] - ro['obs'][i]).max()); rew_err.append(abs(rew_g[i] - ro['reward'][i]))
            flags_ok += int(tg[i] == ro['terminated'][i] and sg[i] == ro['success'][i])
            contacts_ok += int(_pairs_of(po[i], co[i]) == sets_g[i])
        # episodes that ended are restarted on the oracle so that the next round starts from live states
        done = ro['terminated'] | ro['truncated']
        if done.any():
            ob.reset(mask=done)
    ovf = gb.stats()['capacity_overflow_ticks'] if precision else 0
    t_o, t_g = np.array(t_o, float), np.array(t_g, float)
    return dict(env=name, precision=precision or 'oracle+%g' % perturb, cases=total, same_ticks=same_ticks / total,
                obs_err_p50=float(np.median(obs_err)), obs_err_p95=float(np.quantile(obs_err, 0.95)), obs_err_max=float(np.max(obs_err)),
                rew_err_p95=float(np.quantile(rew_err, 0.95)), flags_ok=flags_ok / same_ticks, contacts_ok=contacts_ok / same_ticks,
                ticks_mean_oracle=float(t_o.mean()), ticks_mean_gpu=float(t_g.mean()), tick_diff_p90=float(np.quantile(np.abs(t_o - t_g), 0.9)),
                within_2_ticks=float((np.abs(t_o - t_g) <= 2).mean()), tick_mismatch_margin_p50=float(np.median(tick_margin)) if tick_margin else None,
                tick_mismatch_margin_max=float(np.max(tick_margin)) if tick_margin else None, capacity_overflow_ticks=ovf)


EXACT_SCENES = ('BallBalance', 'TwoBallsBalance', 'CubesPush', 'CubesPushDiff')   # 1e-13 never changes an integer output of the oracle


@pytest.mark.parametrize('name', ENVS)
def test_parity_build_gives_the_oracles_integer_outputs(name):
    """fp64 kernels against the fp64 oracle: trip counts, flags and contact-pair sets EQUAL, or as equal as the oracle is
    to itself under a 1e-13 perturbation (the scenes whose step map is chaotic at round-off level)."""
    m = _resync_run(name, 'f64')
    _record('resync_steps_f64', m)
    if name in EXACT_SCENES:
        # EQUAL, except that a step during which a tick carried more manifold points than the kernel keeps (CubesPush: 24;
        # the checker keeps them all) may end on another tick: never more mismatches than such ticks (measured 1 of 1600)
        mismatched = round((1.0 - m['same_ticks']) * m['cases'])
        assert mismatched <= m['capacity_overflow_ticks'] and mismatched <= 2, m
        assert m['flags_ok'] == 1.0 and m['contacts_ok'] == 1.0, m
        assert m['obs_err_max'] < 1e-5, m          # a 1e-16 difference grows through <= 480 ticks of contact dynamics; typical 1e-13
    else:
        base = _resync_run(name, None, perturb=1e-13)
        _record('resync_steps_oracle_self', base)
        assert m['same_ticks'] >= base['same_ticks'] - 0.03, (m, base)
        assert m['flags_ok'] >= min(base['flags_ok'], 0.995) - 0.005 and m['contacts_ok'] >= base['contacts_ok'] - 0.03, (m, base)
        assert m['obs_err_p50'] < 1e-9, m


@pytest.mark.parametrize('name', ENVS)
def test_resynchronised_steps_match_oracle(name):
    """The product (fp32) against the oracle. The same kernels in fp64 reproduce the oracle (test above), so whatever
    differs here is the scalar width; it is compared with what an fp32-size perturbation does to the oracle itself."""
    m = _resync_run(name, 'f32')
    # the yardstick: fp32 arithmetic moves the state by ~1e-6 in ONE tick (test_single_tick_from_identical_state measures
    # 1e-7..3e-6 on positions), and does so on each of the 10..480 ticks of a step; the oracle under a one-off perturbation of
    # that size is the cheapest honest stand-in for 'the same algorithm in another precision'
    base = _resync_run(name, None, perturb=1e-6)
    m['oracle_self_same_ticks'] = base['same_ticks']
    _record('resync_steps', m); _record('resync_steps_oracle_self', base)
    assert m['same_ticks'] >= base['same_ticks'] - 0.08, (m, base)
    if name in EXACT_SCENES:
        assert m['same_ticks'] >= 0.95 and m['tick_mismatch_margin_max'] is None or m['tick_mismatch_margin_max'] < 3e-3, m
    assert abs(m['ticks_mean_gpu'] / m['ticks_mean_oracle'] - 1) < 0.10, m
    # one bar for every scene (round 1 needed 3e-2 / 90 % for GolfJoints; with the polytope fingers it measures 2.4e-4 / 97.4 %)
    assert m['obs_err_p95'] < 5e-3 and m['flags_ok'] >= 0.99 and m['contacts_ok'] >= 0.95, m
    assert m['rew_err_p95'] < 5e-3, m


# ------------------------------------------------------------------------------- env logic vs the reference fixture
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), 'golden', 'reference_env_logic.json')))
ON_THRESHOLD = ('exactly', '_at_', 'level_with', 'unit_distance')


@pytest.mark.parametrize('name', ENVS)
@pytest.mark.parametrize('reward_type', ['dense', 'sparse'])
def test_env_logic_matches_reference_fixture(name, reward_type):
    from test_reference_env_logic import _oracle_pairs
    if name not in GOLD['envs']:
        pytest.skip('no recorded step() cases for %s (its env logic is checked against the oracle in test_box_env_logic)' % name)
    cases = [c for c in GOLD['envs'][name]['step'] if c['reward_type'] == reward_type and 'raises' not in c
             and not any(t in c['label'] for t in ON_THRESHOLD)]
    if not cases:
        pytest.skip('no %s cases for %s' % (reward_type, name))
    n = len(cases)
    gb = _batch(name, n, seed=1, reward_sparse=(reward_type == 'sparse'))
    gb.reset()
    table = O.OracleBatch(name, 1).shape_table()
    st = np.zeros((n, 96), np.float32); acts = np.zeros((n, gb.act_dim), np.float32); us = np.zeros(n, np.float32)
    pairs = -np.ones((n, 32, 2), np.int32); counts = np.zeros(n, np.int32)     # URGYM_MAX_CONTACTS wide
    for i, c in enumerate(cases):
        st[i, :len(c['state'])] = c['state']; acts[i] = c['action']; us[i] = c['u']
        pr = _oracle_pairs(table, c['contacts'])
        counts[i] = len(pr)
        for k, p in enumerate(pr):
            pairs[i, k] = p
    gb.set_state(torch.tensor(st))
    obs, rew, term, trunc, succ = gb.step_logic(acts, pairs, counts, u=us)
    obs, rew, term, succ = _np(obs), _np(rew), _np(term).astype(bool), _np(succ).astype(bool)
    after = _np(gb.get_state())
    bad = []
    for i, c in enumerate(cases):
        e_obs = np.abs(obs[i] - np.array(c['obs'])).max()
        e_rew = abs(rew[i] - c['reward'])
        ok = e_obs < 2e-5 and e_rew < 2e-4 and term[i] == c['terminated'] and succ[i] == bool(c['info'].get('is_success', False))
        cmd = c['commands'][0]
        if cmd['op'] == 'ik':      # the commanded joint targets are the IK solution for the reference's target pose
            want = O.ik(st[i, :12].astype(np.float64), np.array(cmd['pos'] + cmd['orn']))
            ok = ok and np.abs(after[i, 24:30] - want).max() < 1e-4
        elif cmd['op'] == 'joints':   # CubesGrasp: the action itself (exact); GolfJoints: q + action, summed in fp32 here
            ok = ok and np.abs(after[i, 24:30] - np.array(cmd['targets'])).max() <= (0.0 if name == 'CubesGrasp' else 1e-6)
        else:
            ok = ok and abs(after[i, 30] - cmd['target']) < 1e-6
        if name == 'CubesGrasp':
            fl = int(after[i, 31])
            ok = ok and {'approached': bool(fl & 4), 'grasped': bool(fl & 8)} == c['subgoals_after']
        if not ok:
            bad.append((i, c['label'], c['contacts'], float(e_obs), float(e_rew), bool(term[i]), c['terminated']))
    _record('env_logic', dict(env=name, reward_type=reward_type, cases=n, bad=len(bad)))
    assert not bad, bad[:5]


# ------------------------------------------------------------------------------------------ full-size properties
@pytest.mark.parametrize('name,n', [('BallBalance', 4096), ('TwoBallsBalance', 16384), ('CubesPush', 16384), ('CubesGrasp', 16384), ('Golf', 4096), ('Box', 4096)])
def test_full_size_run_properties(name, n):
    gb = _batch(name, n, seed=11)
    gb.reset()
    assert int(_np(gb.reset_ticks).min()) >= (10 if name == 'Box' else 11)
    rng = np.random.default_rng(3)
    steps = 22 if name == 'CubesGrasp' else 8
    ended = 0; tick_sum = int(_np(gb.reset_ticks).astype(np.int64).sum())
    restarted_early = torch.zeros(n, dtype=torch.bool, device=gb.device)
    for k in range(steps):
        a = torch.tensor(_actions(rng, name, n), dtype=torch.float32, device=gb.device)
        obs, rew, term, trunc, succ = gb.step(a, autoreset=True)
        assert torch.isfinite(obs).all() and torch.isfinite(rew).all()
        done = (term | trunc).bool()
        ended += int(done.sum()); tick_sum += int(gb.nticks.to(torch.int64).sum())
        assert not (succ.bool() & ~term.bool()).any()                                   # success implies termination
        nt = gb.nticks[:, 0]
        assert int(nt.min()) >= 1 and int(nt.max()) <= 480
        assert ((gb.nticks[:, 1] > 0) == done).all()                                    # reset ticks only where an episode ended
        if name == 'CubesGrasp' and k < 19:
            restarted_early |= done
        if name == 'CubesGrasp' and k == 19:
            assert trunc.bool()[~restarted_early].all()                                 # TimeLimit: 20 steps
    st = gb.get_state()
    assert torch.isfinite(st).all()
    if name == 'Box':                                                                   # anchored bodies stay on their joints and within their limits
        assert (gb.nticks[:, 0] == 120).all()
        assert (st[:, 38:41] - torch.tensor([0.075, 0.0, 0.1], device=st.device)).abs().max() < 1e-6 and st[:, 51:53].abs().max() < 1e-6
        door = 2 * torch.atan2(st[:, 42], st[:, 44]); button = st[:, 53] - 0.07
        assert st[:, [41, 43]].abs().max() < 1e-6 and door.min() > -0.02 and door.max() < 2.02 and button.min() > -0.052 and button.max() < 2e-3
    fl = st[:, 31].to(torch.int64)
    for b in range(3):
        act = (fl & (16 << b)) != 0
        if act.any():
            qn = st[act][:, 41 + 13 * b: 45 + 13 * b].norm(dim=1)
            assert (qn - 1).abs().max() < 1e-4
    assert (st[:, 32] >= 0).all() and (st[:, 32] < gb.max_episode_steps).all()
    stats = gb.stats()
    assert stats['episodes'] == ended and stats['env_steps'] == steps * n and stats['ticks'] == tick_sum
    assert stats['length_sum'] >= stats['episodes']


@pytest.mark.parametrize('name,n,steps,bar', [('BallBalance', 4096, 12, 0.0), ('TwoBallsBalance', 16384, 6, 0.0), ('CubesPush', 16384, 12, 3e-4),
                                              ('CubesGrasp', 16384, 4, 1e-4), ('Golf', 4096, 6, 0.02)])
def test_point_capacity_rarely_binds_on_the_benchmarked_workloads(name, n, steps, bar):
    """The kernel keeps 20 manifold points per tick (24 in the CubesPush scenes, 32 in the Golf scenes); the oracle keeps all. On
    the BASELINE configs, at their full batch sizes and under the benchmarked random policy, the fraction of ticks that had more
    (urgym_stats out[6]) must be zero (ball scenes), below 1e-4 (CubesGrasp; measured 5e-6), below 3e-4 (CubesPush; measured
    1.2e-4: the random policy drives both fingers into the table next to the cubes) and below 2 % for Golf (measured 0.4 %).
    The contact predicates never depend on the capacity (per-pair bits are set before the cut)."""
    gb = _batch(name, n, seed=17)
    gb.reset()
    rng = np.random.default_rng(8)
    for _ in range(steps):
        gb.step(torch.tensor(_actions(rng, name, n), dtype=torch.float32, device=gb.device), autoreset=True)
    st = gb.stats()
    frac = st['capacity_overflow_ticks'] / st['ticks']
    _record('capacity', dict(env=name, envs=n, ticks=st['ticks'], overflow_ticks=st['capacity_overflow_ticks'], fraction=frac))
    assert frac <= bar, (name, frac)


@pytest.mark.parametrize('name', ['BallBalance', 'CubesPush'])
def test_shard_layout_is_bit_exact(name):
    whole, part = _batch(name, 64, seed=13), _batch(name, 32, seed=13, env_index_offset=32)
    np.testing.assert_array_equal(_np(whole.reset())[32:], _np(part.reset()))
    rng = np.random.default_rng(4)
    for _ in range(3):
        a = _actions(rng, name, 64)
        whole.step(a, autoreset=True); part.step(a[32:], autoreset=True)
        np.testing.assert_array_equal(_np(whole.obs)[32:], _np(part.obs))
        np.testing.assert_array_equal(_np(whole.nticks)[32:], _np(part.nticks))
    np.testing.assert_array_equal(_np(whole.get_state())[32:], _np(part.get_state()))


def test_host_entry_equals_device_entry():
    import ctypes as C
    from urgym_b200 import _lib
    n = 48
    a_dev, a_host = _batch('CubesPush', n, seed=2), _batch('CubesPush', n, seed=2)
    a_dev.reset(); a_host.reset()
    L = _lib.lib()
    rng = np.random.default_rng(6)
    for _ in range(3):
        act = _actions(rng, 'CubesPush', n).astype(np.float32)
        a_dev.step(act, autoreset=True)
        obs = np.zeros((n, 9), np.float32); rew = np.zeros(n, np.float32); fin = np.zeros((n, 9), np.float32)
        t, tr, su = (np.zeros(n, np.uint8) for _ in range(3)); nt = np.zeros((n, 2), np.int32)
        p = lambda x: x.ctypes.data_as(C.c_void_p)
        _lib.check(L.urgym_step_host(a_host.h, p(act), None, None, p(obs), p(rew), p(t), p(tr), p(su), p(fin), p(nt), 1))
        np.testing.assert_array_equal(obs, _np(a_dev.obs)); np.testing.assert_array_equal(rew, _np(a_dev.reward))
        np.testing.assert_array_equal(nt, _np(a_dev.nticks)); np.testing.assert_array_equal(t, _np(a_dev.terminated))


def test_state_round_trip_and_masked_reset():
    gb = _batch('TwoBallsBalance', 16, seed=8)
    gb.reset()
    s = gb.get_state().clone()
    gb.set_state(s)
    assert torch.equal(gb.get_state(), s)
    obs_before = gb.obs.clone()
    mask = torch.zeros(16, dtype=torch.uint8); mask[::4] = 1
    gb.reset(mask=mask)
    s2 = gb.get_state()
    keep = ~mask.bool().to(s.device)
    assert torch.equal(s2[keep], s[keep]) and torch.equal(gb.obs[keep], obs_before[keep])
    assert (s2[~keep][:, 33] == s[~keep][:, 33] + 1).all()                              # a new episode where asked


def test_determinism_across_runs():
    outs = []
    for _ in range(2):
        gb = _batch('Golf', 64, seed=21)
        gb.reset()
        rng = np.random.default_rng(9)
        for _ in range(3):
            gb.step(_actions(rng, 'Golf', 64), autoreset=True)
        outs.append((_np(gb.obs).copy(), _np(gb.get_state()).copy()))
    np.testing.assert_array_equal(outs[0][0], outs[1][0]); np.testing.assert_array_equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize('name', ['BallBalance', 'Golf'])
def test_reset_launch_scheduling_leaves_results_unchanged(name, monkeypatch):
    """Auto-resets run inside the step launch or in a reset-only launch that follows it; which one is a
    scheduling choice per env type and must not change a single bit of what the caller sees."""
    outs = []
    for split in ('0', '1'):
        monkeypatch.setenv('URGYM_SPLIT_RESET', split)
        gb = _batch(name, 3000, seed=23)
        gb.reset()
        rng = np.random.default_rng(12)
        rec = []
        for _ in range(6):
            gb.step(_actions(rng, name, 3000), autoreset=True)
            rec.append([_np(x).copy() for x in (gb.obs, gb.reward, gb.terminated, gb.truncated, gb.success, gb.final_obs, gb.nticks)])
        rec.append([_np(gb.get_state()).copy()])
        outs.append(rec)
        assert gb.stats()['episodes'] > 0
    for a, b in zip(outs[0], outs[1]):
        for x, y in zip(a, b):
            np.testing.assert_array_equal(x, y)


@pytest.mark.parametrize('name', ['BallBalance', 'TwoBallsBalance', 'CubesGrasp'])
def test_register_resident_sweeps_leave_results_unchanged(name, monkeypatch):
    """Contact-free ticks with 12 or 13 joint rows run their PGS sweeps out of registers (dev_solver.cuh,
    joint_sweeps_reg): same operations in the same order as the generic sweeps, so not a bit may differ."""
    outs = []
    for generic in ('1', '0'):
        monkeypatch.setenv('URGYM_GENERIC_SWEEPS', generic)
        gb = _batch(name, 1024, seed=29)
        gb.reset()
        rng = np.random.default_rng(14)
        rec = [[_np(gb.obs).copy(), _np(gb.reset_ticks).copy()]]
        for _ in range(6):
            gb.step(_actions(rng, name, 1024), autoreset=True)
            rec.append([_np(x).copy() for x in (gb.obs, gb.reward, gb.terminated, gb.truncated, gb.success, gb.final_obs, gb.nticks)])
        rec.append([_np(gb.get_state()).copy()])
        outs.append(rec)
    for a, b in zip(outs[0], outs[1]):
        for x, y in zip(a, b):
            np.testing.assert_array_equal(x, y)


@pytest.mark.parametrize('name', ['BallBalance', 'CubesGrasp', 'Golf'])
def test_distributional_agreement_over_episodes(name):
    """Free-running (no re-synchronisation) rollouts diverge trajectory by trajectory — the dynamics are chaotic — but
    the statistics a learner sees must agree: mean settle ticks per step within 5% (Golf 10%), termination rate within 0.05."""
    from concurrent.futures import ThreadPoolExecutor
    n, steps = 256, 24
    ob, gb = O.OracleBatch(name, n, seed=31), _batch(name, n, seed=31)
    rng = np.random.default_rng(10)
    acts = [_actions(rng, name, n) for _ in range(steps)]
    us = [rng.uniform(0, 1, n).astype(np.float32).astype(np.float64) for _ in range(steps)]
    gb.reset()
    nthr = min(16, os.cpu_count() or 1)
    cuts = np.linspace(0, n, nthr + 1).astype(int)
    with ThreadPoolExecutor(nthr) as ex:
        list(ex.map(lambda i: ob.reset(lo=int(cuts[i]), hi=int(cuts[i + 1])), range(nthr)))
        t_o = t_g = d_o = d_g = 0
        for k in range(steps):
            res = list(ex.map(lambda i: ob.step(acts[k], u=us[k], autoreset=True, lo=int(cuts[i]), hi=int(cuts[i + 1])), range(nthr)))
            for i, r in enumerate(res):
                sl = slice(int(cuts[i]), int(cuts[i + 1]))
                t_o += int(r['nticks'][sl, 0].sum()); d_o += int((r['terminated'][sl] | r['truncated'][sl]).sum())
            gb.step(acts[k], u=us[k], autoreset=True)
            t_g += int(gb.nticks[:, 0].sum()); d_g += int((gb.terminated | gb.truncated).bool().sum())
    m = dict(env=name, ticks_per_step_oracle=t_o / (n * steps), ticks_per_step_gpu=t_g / (n * steps), done_rate_oracle=d_o / (n * steps), done_rate_gpu=d_g / (n * steps))
    _record('distribution', m)
    assert abs(m['ticks_per_step_gpu'] / m['ticks_per_step_oracle'] - 1) < (0.10 if name == 'Golf' else 0.05), m
    assert abs(m['done_rate_gpu'] - m['done_rate_oracle']) < 0.05, m


EPISODE_STATS = os.path.join(os.path.dirname(__file__), 'golden', 'episode_stats_oracle.json')


@pytest.mark.parametrize('name,n,steps', [('BallBalance', 4096, 56), ('CubesPush', 1024, 500), ('CubesGrasp', 1024, 200)])
def test_episode_statistics_over_1e4_episodes(name, n, steps):
    """>= 1e4 finished episodes per side under the same seeded random policy: mean return, mean length and success rate of the
    CUDA path agree with the oracle's within 4 standard errors (plus 1% of the value for the means). The oracle's side (40 M
    ticks) is recorded once by tests/golden/make_episode_stats.py with the same seeds, env counts and step counts
    (tests/golden/episode_stats_oracle.json); only the device rolls out here."""
    want = json.load(open(EPISODE_STATS))[name]
    assert (want['envs'], want['steps'], want['seed_env'], want['seed_actions']) == (n, steps, 41, 11), want
    gb = _batch(name, n, seed=41)
    rng = np.random.default_rng(11)
    gb.reset()
    ret_g, len_g = np.zeros(n), np.zeros(n, int)
    eps = []
    for k in range(steps):
        a = _actions(rng, name, n); u = rng.uniform(0, 1, n).astype(np.float32).astype(np.float64)
        gb.step(a, u=u, autoreset=True)
        rew_g, done_g, suc_g = _np(gb.reward), _np((gb.terminated | gb.truncated).bool()), _np(gb.success.bool())
        ret_g += rew_g; len_g += 1
        for i in np.nonzero(done_g)[0]:
            eps.append((ret_g[i], len_g[i], suc_g[i]))
        ret_g[done_g] = 0; len_g[done_g] = 0
    eg = np.array(eps, float)
    m = dict(env=name, episodes_oracle=want['episodes'], episodes_gpu=len(eg))
    assert want['episodes'] >= 10000 and len(eg) >= 10000, m
    for j, key in enumerate(('return', 'length', 'success')):
        mo, mg = want['mean'][key], eg[:, j].mean()
        se = np.sqrt(want['var'][key] / want['episodes'] + eg[:, j].var() / len(eg))
        m['%s_oracle' % key], m['%s_gpu' % key], m['%s_se' % key] = float(mo), float(mg), float(se)
    _record('episode_statistics', m)
    for key in ('return', 'length', 'success'):
        tol = 4 * m['%s_se' % key] + (0.01 * abs(m['%s_oracle' % key]) if key != 'success' else 0.002)
        assert abs(m['%s_gpu' % key] - m['%s_oracle' % key]) <= tol, (key, m)
    st = gb.stats()                                                             # the device-side counters tell the same story
    assert st['episodes'] == len(eg) and st['successes'] == eg[:, 2].sum()
    assert st['return_sum'] == pytest.approx(eg[:, 0].sum(), rel=1e-4) and st['length_sum'] == eg[:, 1].sum()


# ------------------------------------------------------------------------------------------- a grasp that happens
@pytest.mark.parametrize('precision', ['f64', 'f32'])
def test_scripted_grasp_sub_goals_and_success_match_the_oracle(precision):
    """CubesGrasp's hardest contact path driven through the physics on both sides (tests/scripted_grasp.py): open, move above
    the cube, down around it, close, raise. Free-running (no re-synchronisation) from the same reset with the same actions:
    the sticky sub-goal bits, terminated and is_success must be equal at every step, rewards to 1e-3 (fp32: 2e-2 on the dense
    distance term before 'approached'), and the cube must end above 0.2 m in the gripper on both sides."""
    import scripted_grasp as G
    n = 8
    ob, gb = O.OracleBatch('CubesGrasp', n, seed=11), _batch('CubesGrasp', n, seed=11, precision=precision)
    ob.reset(); gb.reset()
    rows = []

    def on_step(k, a, ro):
        gb.step(a, u=np.full(n, 0.5), autoreset=False)
        rows.append(dict(bits=_np(gb.obs)[:, :2].copy(), rew=_np(gb.reward).astype(np.float64).copy(), term=_np(gb.terminated).astype(bool).copy(),
                         succ=_np(gb.success).astype(bool).copy(), cube_z=_np(gb.obs)[:, 12].copy(), o_bits=ro['obs'][:, :2].copy(), o_rew=ro['reward'].copy(),
                         o_term=ro['terminated'].copy(), o_succ=ro['success'].copy()))
    log = G.run(ob, n, on_step)
    for k, r in enumerate(rows):
        assert np.array_equal(r['bits'], r['o_bits']) and np.array_equal(r['term'], r['o_term']) and np.array_equal(r['succ'], r['o_succ']), (k, r)
        assert np.abs(r['rew'] - r['o_rew']).max() < (1e-3 if precision == 'f64' else 2e-2), (k, r)
    assert rows[-1]['succ'].all() and (rows[-1]['cube_z'] > 0.2).all() and (rows[-1]['bits'] == 1).all()
    _record('scripted_grasp', dict(precision=precision, steps=len(rows), success=float(rows[-1]['succ'].mean()), rew_err_max=float(max(np.abs(r['rew'] - r['o_rew']).max() for r in rows))))


# ----------------------------------------------------------------------------------- BASELINE config 0, shipped actors
@pytest.mark.parametrize('precision', ['f64', 'f32'])
def test_config0_correctness_trace_replays_on_the_device(precision):
    """BASELINE.json config 0: BallBalance, one env, 1000 gym steps with auto-reset, actions from torch.Generator(0), reset
    draws from random.Random(0) (tests/golden/make_config0_trace.py, recorded on the fp64 oracle). Every step is replayed from
    the recorded state: the fp64 build reproduces every integer of it, the fp32 product all but a few threshold flips."""
    from urgym_b200 import trace as T
    tr = T.load(os.path.join(os.path.dirname(__file__), 'golden', 'config0_ballbalance_trace.npz'))
    gb = _batch('BallBalance', 1, seed=0, precision=precision); gb.reset()
    d = T.replay(gb, tr, resync=True)
    m = dict(precision=precision, steps=len(d['same_ticks']), same_ticks=float(d['same_ticks'].mean()), same_flags=float(d['same_flags'].mean()),
             same_contacts=float(d['same_contacts'].mean()), obs_err_p50=float(np.median(d['obs_err'])), obs_err_max=float(d['obs_err'].max()))
    _record('config0_trace', m)
    if precision == 'f64':
        assert m['same_ticks'] == 1.0 and m['same_flags'] == 1.0 and m['same_contacts'] == 1.0 and m['obs_err_max'] < 1e-6, m
    else:
        assert m['same_ticks'] >= 0.95 and m['same_flags'] >= 0.995 and m['same_contacts'] >= 0.95 and m['obs_err_p50'] < 1e-4, m


ORACLE_REPLAY = os.path.join(os.path.dirname(__file__), 'golden', 'actor_replay_oracle_1e4.json')


@pytest.mark.parametrize('name', ['TwoBallsBalance', 'CubesPush', 'Golf', 'GolfJoints'])
def test_shipped_actors_on_the_device_match_the_oracle_replay(name):
    """The reference's four shipped actors (pybullet-trained; tests/golden/actors.npz) drive 1e4 episodes on the CUDA backend
    with the MLP on the device (tools/replay_actors.py --backend cuda). Return, length and success rate must agree with the
    same replay on the fp64 oracle (1e4 episodes, recorded in tests/golden/actor_replay_oracle_1e4.json by
    `tools/replay_actors.py --episodes 10000`) within 4 standard errors plus 2 % of the value. Both are printed next to the
    checkpoints' pybullet statistics in profiles/."""
    sys_path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tools')
    import sys
    sys.path.insert(0, sys_path)
    import replay_actors as R
    want = json.load(open(ORACLE_REPLAY))[name]['replay']
    got = R.rollout_cuda(name, 10000, seed=123)
    m = dict(env=name, device=got, oracle=want, pybullet=json.load(open(ORACLE_REPLAY))[name]['pybullet_training_last_100'])
    _record('actor_replay', m)
    # ticks that carried more manifold points than the kernel keeps, under the trained policies (which work with the fingers on
    # the table, next to the cubes): measured 0.2 % of the ticks for CubesPush, below 2 % everywhere
    assert got['capacity_overflow_ticks'] / max(got['ticks'], 1) < 0.02, m
    for key in ('return', 'length', 'success'):
        a, b_ = got['mean_%s' % key if key != 'success' else 'success_rate'], want['mean_%s' % key if key != 'success' else 'success_rate']
        se = np.hypot(got['se_%s' % key], want['se_%s' % key])
        assert abs(a - b_) <= 4 * se + 0.02 * abs(b_) + (0.01 if key == 'success' else 0.0), (key, m)


# ------------------------------------------------------------------------------------------------------- Box-v0
@pytest.mark.parametrize('precision', ['f64', 'f32'])
def test_box_env_logic(precision):
    """Box-v0's step() around the physics (env_box_v0.py:124-159), device against oracle through the logic-only entries
    (urgym_step_logic / the oracle's prologue + epilogue): the commanded IK target is the end-effector pose plus the first seven
    action components, the gripper target comes from the eighth, and the three sticky sub-goals fire in order from states placed
    by hand — door beyond 1.9 rad (opened), button below -0.02 m (pressed; any toucher: button_touch_mode='any'), door back under
    0.1 rad (closed -> terminated, success); reward +1 - 0.1 at each transition, -0.1 otherwise."""
    ob, gb = O.OracleBatch('Box', 1, seed=2), _batch('Box', 1, seed=2, precision=precision)
    ob.reset(); gb.reset()
    npdt = np.float64 if precision == 'f64' else np.float32
    rng = np.random.default_rng(3)
    none = -np.ones((1, 32, 2), np.int32); zero = np.zeros(1, np.int32)
    script = [(0.0, 0.0, -0.1, False), (1.85, 0.0, -0.1, False), (1.95, 0.0, 0.9, False), (1.95, -0.01, -0.1, False), (0.05, -0.03, 0.9, False),
              (1.0, 0.0, -0.1, False), (0.05, 0.0, 0.9, True)]
    for k, (door, button, rew, term) in enumerate(script):
        s = ob.get_state()
        s[0, 41:45] = [0, np.sin(door / 2), 0, np.cos(door / 2)]; s[0, 53] = 0.07 + button
        s = s.astype(npdt)
        ob.set_state(s.astype(np.float64)); gb.set_state(torch.tensor(s))
        a = _actions(rng, 'Box', 1)
        pro = ob.prologue(a[0])
        ro = ob.epilogue(np.zeros((0, 2), np.int32), reward_so_far=pro['reward'])
        obs, r_g, t_g, _, s_g = gb.step_logic(a, none, zero)
        after = _np(gb.get_state()).astype(np.float64)
        assert ro['reward'] == pytest.approx(rew, abs=1e-12) and ro['terminated'] == term and ro['success'] == term, (k, ro)
        assert abs(float(r_g[0]) - rew) < 1e-6 and bool(t_g[0]) == term and bool(s_g[0]) == term, (k, float(r_g[0]), bool(t_g[0]))
        tol = 1e-9 if precision == 'f64' else 2e-5
        assert np.abs(_np(obs)[0].astype(np.float64) - ro['obs']).max() < tol, k
        assert np.abs(after[0, 24:31] - pro['targets']).max() < (1e-9 if precision == 'f64' else 1e-4), (k, after[0, 24:31], pro['targets'])
        fl = int(after[0, 31])
        assert bool(fl & 4) == (k >= 2) and bool(fl & 8) == (k >= 4), (k, fl)


@pytest.mark.parametrize('precision', ['f64', 'f32'])
def test_box_button_press_through_the_physics(precision):
    """Box-v0's new physics (world-anchored door and button, their limit and motor rows, finger-on-button contacts) driven on both
    sides, free-running from the same state with the same actions: the door stands open on its upper limit, the closed gripper
    comes down on the button and pushes it past -0.02 m against its 1 N motor: 'button_pressed' fires on the same step with the
    same reward on the device and on the oracle, the door never leaves its limit, the button ends within its stroke."""
    n = 4
    ob, gb = O.OracleBatch('Box', n, seed=1), _batch('Box', n, seed=1, precision=precision)
    ob.reset(); gb.reset()
    npdt = np.float64 if precision == 'f64' else np.float32
    s = ob.get_state(); s[:, 41:45] = [0, np.sin(1.0), 0, np.cos(1.0)]; s[:, 31] = s[:, 31].astype(int) | 4
    s = s.astype(npdt); ob.set_state(s.astype(np.float64)); gb.set_state(torch.tensor(s))
    ro = ob.step(np.zeros((n, 8))); gb.step(np.zeros((n, 8)), autoreset=False)
    pressed_at = None
    for k in range(12):
        d = np.clip(np.array([0.0, 0.0, 0.05]) - ro['obs'][:, :3], -0.05, 0.05)
        a = np.zeros((n, 8)); a[:, :3] = d
        a = a.astype(np.float32).astype(np.float64)
        ro = ob.step(a); gb.step(a, autoreset=False)
        fo, fg = ob.get_state()[:, 31].astype(int) & 12, _np(gb.get_state())[:, 31].astype(int) & 12
        assert np.array_equal(fo, fg), (k, fo, fg)
        assert np.abs(_np(gb.reward).astype(np.float64) - ro['reward']).max() < 1e-6, (k, _np(gb.reward), ro['reward'])
        assert np.abs(_np(gb.obs).astype(np.float64) - ro['obs']).max() < (1e-7 if precision == 'f64' else 5e-3), k
        if (fo & 8).all():
            pressed_at = k
            break
    assert pressed_at is not None and np.allclose(ro['reward'], 0.9)
    sg = _np(gb.get_state()).astype(np.float64)
    door = 2 * np.arctan2(sg[:, 42], sg[:, 44]); button = sg[:, 53] - 0.07
    assert np.abs(door - 2.0).max() < 5e-3 and (button < -0.02).all() and (button > -0.0505).all(), (door, button)
    _record('box_button_press', dict(precision=precision, pressed_at_step=int(pressed_at), button=float(button.mean())))


@pytest.mark.parametrize('precision', ['f64', 'f32'])
def test_box_task_solved_by_the_scripted_policy_on_the_device(precision):
    """The whole Box-v0 task through the physics on the device (tests/scripted_box.py: grasp the handle bar, lift the door along
    its arc past 1.9 rad, let go, press the button, push the door shut): the three +1 transitions and the final terminated /
    is_success happen on the device as on the oracle — on the same steps in fp64, within 2 steps in fp32 (free-running, ~17 k
    ticks of contact dynamics per env)."""
    import scripted_box as SB
    n = 2
    ob, gb = O.OracleBatch('Box', n, seed=1), _batch('Box', n, seed=1, precision=precision)
    ob.reset(); gb.reset()
    log_o = SB.run(lambda a: ob.step(a), n)

    def dev_step(a):
        gb.step(a, autoreset=False)
        return dict(obs=_np(gb.obs).astype(np.float64), reward=_np(gb.reward).astype(np.float64), terminated=_np(gb.terminated).astype(bool),
                    success=_np(gb.success).astype(bool))
    log_g = SB.run(dev_step, n)
    wo = [k for k, r in enumerate(log_o) if (r['reward'] > 0.5).all()]
    wg = [k for k, r in enumerate(log_g) if (r['reward'] > 0.5).all()]
    _record('box_scripted_task', dict(precision=precision, steps_oracle=len(log_o), steps_device=len(log_g), wins_oracle=wo, wins_device=wg))
    assert len(wo) == 3 and len(wg) == 3, (wo, wg)
    assert log_g[-1]['terminated'].all() and log_g[-1]['success'].all() and len(log_g) < 200
    tol = 0 if precision == 'f64' else 2
    assert all(abs(a - b) <= tol for a, b in zip(wo, wg)) and abs(len(log_o) - len(log_g)) <= tol, (wo, wg)
    assert gb.stats()['capacity_overflow_ticks'] <= 0.01 * gb.stats()['ticks']


def test_box_step_logic_matches_the_reference_fixture():
    """The device env logic of Box-v0 against the cases recorded from the reference's own BoxManipulation.step()
    (tests/golden/reference_box_logic.json): observation <= 2e-5, reward <= 2e-6, terminated / is_success / sticky bits exact,
    commanded joint targets = the IK solution for the reference's target pose (position + quaternion, each incremented
    component-wise) and its gripper target. Cases that sit exactly on a threshold are excluded (fp32 may land on either side)."""
    G = json.load(open(os.path.join(os.path.dirname(__file__), 'golden', 'reference_box_logic.json')))
    cases = [c for c in G['step'] if 'exactly' not in c['label']]
    n = len(cases)
    gb = _batch('Box', n, seed=1); gb.reset()
    st = np.zeros((n, 96), np.float32); acts = np.zeros((n, 8), np.float32)
    for i, c in enumerate(cases):
        st[i] = c['state']; acts[i] = c['action']
    gb.set_state(torch.tensor(st))
    obs, rew, term, _, succ = gb.step_logic(acts, -np.ones((n, 32, 2), np.int32), np.zeros(n, np.int32))
    after = _np(gb.get_state())
    for i, c in enumerate(cases):
        assert np.abs(_np(obs)[i] - np.array(c['obs'])).max() < 2e-5 and abs(float(rew[i]) - c['reward']) < 2e-6, (c['label'], float(rew[i]), c['reward'])
        assert bool(term[i]) == c['terminated'] and bool(succ[i]) == c['info']['is_success'], c['label']
        fl = int(after[i, 31])
        assert bool(fl & 4) == c['flags_after']['box_opened'] and bool(fl & 8) == c['flags_after']['button_pressed'], c['label']
        ik = [x for x in c['commands'] if x['op'] == 'ik'][0]; grip = [x for x in c['commands'] if x['op'] == 'gripper'][0]
        want = O.ik(st[i, :12].astype(np.float64), np.array(ik['pos'] + ik['orn']), base=(0, 0.5, 0))
        assert np.abs(after[i, 24:30] - want).max() < 1e-4 and abs(after[i, 30] - grip['target']) < 1e-6, c['label']
    _record('env_logic', dict(env='Box', reward_type='dense', cases=n, bad=0))


# ------------------------------------------------------------------------------------------------- gym-style API
def test_vector_env_and_single_env_api():
    from urgym_b200 import envs as E
    v = E.URGymVectorEnv('URGym/CubesGrasp-v0', 128, seed=1)
    obs, info = v.reset()
    assert obs.shape == (128, 17) and obs.is_cuda and 'reset_ticks' in info
    for k in range(20):
        a = torch.rand(128, 9, device=obs.device)
        obs, rew, term, trunc, info = v.step(a)
    assert trunc.dtype == torch.bool and (term | trunc).all()                       # TimeLimit of 20 reached on every env
    assert info['final_observation'].shape == obs.shape and info['is_success'].dtype == torch.bool
    st = v.episode_statistics()
    assert st['episodes'] >= 128 and st['env_steps'] == 20 * 128
    v.close()
    e = E.make('URGym/CubesPush-v0', render_mode='train')
    o, inf = e.reset()
    assert o.dtype == np.float64 and o.shape == (9,) and inf == {} and e.observation_space.contains(np.clip(o, -1, 1))
    for k in range(50):
        o, r, term, trunc, inf = e.step(e.action_space.sample())
        assert isinstance(r, float) and isinstance(term, bool) and 'is_success' in inf
        if term or trunc:
            break
    assert k == 49 and trunc or term
    e.close()
    # Box-v0 through the reference's constructor signature (env_box_v0.py:21) and through the vector env
    e = E.make('URGym/Box-v0', button_touch_mode='any', render_mode='train')
    o, inf = e.reset()
    assert o.shape == (7,) and o.dtype == np.float64 and abs(np.linalg.norm(o[3:]) - 1) < 1e-5 and e.action_space.shape == (8,)
    o, r, term, trunc, inf = e.step(e.action_space.sample())
    assert r == pytest.approx(-0.1, abs=1e-6) and not term and not trunc and inf == {'is_success': False}
    e.close()
    v = E.URGymVectorEnv('URGym/Box-v0', 64, seed=2)
    obs, info = v.reset()
    assert obs.shape == (64, 7) and (info['reset_ticks'] == 10).all()
    obs, rew, term, trunc, info = v.step(torch.zeros(64, 8, device=obs.device))
    assert (info['nticks'][:, 0] == 120).all() and not term.any() and v.max_episode_steps == 200
    v.close()


def test_sb3_vec_env_protocol():
    """SB3 VecEnv conventions (urgym_b200/sb3_vec_env.py): numpy float32 in / out, dones = terminated | truncated with the
    next episode's first observation returned, terminal_observation / TimeLimit.truncated / is_success in infos."""
    from urgym_b200.sb3_vec_env import URGymSB3VecEnv
    v = URGymSB3VecEnv('URGym/CubesGrasp-v0', 64, seed=3)
    obs = v.reset()
    assert obs.shape == (64, 17) and obs.dtype == np.float32
    ref = _batch('CubesGrasp', 64, seed=3); ref.reset()
    rng = np.random.default_rng(12)
    for k in range(21):
        a = _actions(rng, 'CubesGrasp', 64).astype(np.float32)
        obs, rew, dones, infos = v.step(a)
        ref.step(a, autoreset=True)
        np.testing.assert_array_equal(obs, _np(ref.obs))                       # same kernel, same seeds: bit-exact
        np.testing.assert_array_equal(rew, _np(ref.reward))
        assert dones.dtype == bool and len(infos) == 64 and all('is_success' in i for i in infos)
        if k == 19:
            assert dones.all()
            assert all(i['TimeLimit.truncated'] and i['terminal_observation'].shape == (17,) for i in infos)
            np.testing.assert_array_equal(np.stack([i['terminal_observation'] for i in infos]), _np(ref.final_obs))
        elif k < 19:
            assert not any('terminal_observation' in i for i, d in zip(infos, dones) if not d)
    assert v.env_is_wrapped(object) == [False] * 64 and v.get_attr('num_envs', [0, 1]) == [64, 64]
    v.close()


def test_trace_recorded_on_the_device_replays_on_the_oracle(tmp_path):
    from urgym_b200 import trace as T
    gb = _batch('CubesPush', 16, seed=6); gb.reset()
    rng = np.random.default_rng(13)
    tr = T.record(gb, 'CubesPush', rng.uniform(-0.1, 0.1, (6, 16, 3)).astype(np.float32))
    T.save(str(tmp_path / 't.npz'), tr)
    ob = O.OracleBatch('CubesPush', 16, seed=1); ob.reset()
    d = T.replay(ob, T.load(str(tmp_path / 't.npz')), resync=True)
    assert d['same_ticks'].mean() >= 0.9 and d['same_flags'].min() == 1.0 and np.median(d['obs_err']) < 1e-4, d
