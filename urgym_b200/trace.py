"""Rollout traces: a stable on-disk format for "identical initial states and action sequences" comparisons
(SURVEY.md §8f rank 3).

A trace holds everything needed to re-run a rollout on another backend and everything needed to compare the result:
the state records before the first step, the actions (and CubesGrasp selector uniforms), and per step the observation,
reward, flags, settle-loop trip counts, contact-point lists (shape-index pairs, one per manifold point, -1 padded to 64) and
state records. It is written as a NumPy `.npz`.

Any backend exposing `set_state(state[N,96])`, `get_state()`, `step(actions, u=..., autoreset=False)` and
`contacts()` can record or replay: `urgym_b200.batch.EnvBatch` (CUDA), the CPU checker's ctypes binding under `tests/`,
and — on a machine that has pybullet — a recorder that fills the same record from `getJointState` /
`getBasePositionAndOrientation` / `getBaseVelocity`.

State record, 96 float words per env (integers stored as exact floats):
    0..11   joint positions of the 12 movable joints, Bullet joint ids [1,2,3,4,5,6,9,11,13,14,16,18]
    12..23  joint velocities
    24..30  position-motor targets: six arm joints, finger_joint
    31      flags: 1 arm motors armed, 2 gripper motor armed, 4 'approached', 8 'grasped', 16<<b free body b present
    32, 33  step counter of the episode, episode number
    34..37  vertical_quaternion (CubesPush / Golf), xyzw
    38+13b  free body b (b = 0..2): position 3, quaternion xyzw 4, linear velocity 3, angular velocity 3
    77      running episode return (device backend only)
"""
import numpy as np

FORMAT_VERSION = 2          # 2: `pairs` is padded to PAIRS_WIDTH columns whatever the backend's export width
PAIRS_WIDTH = 64
FIELDS = ('state0', 'actions', 'u', 'obs', 'reward', 'terminated', 'truncated', 'success', 'nticks', 'pairs', 'counts', 'state')


def _np(x):
    if hasattr(x, 'detach'):
        return x.detach().cpu().numpy()
    return np.asarray(x)


def _step(backend, a, u, draws=None, autoreset=False):
    out = backend.step(a, u=u, draws=draws, autoreset=autoreset)
    if isinstance(out, dict):          # the CPU checker's binding returns a dict
        return out['obs'], out['reward'], out['terminated'], out['truncated'], out['success'], out['nticks']
    obs, rew, term, trunc, succ = out  # EnvBatch
    return _np(obs), _np(rew), _np(term), _np(trunc), _np(succ), _np(backend.nticks)


def record(backend, env_name, actions, u=None, draws=None):
    """Run `actions[T,N,A]` from the backend's current state and return the trace dict. Without `draws` no env is reset
    (a finished env keeps stepping). With `draws[T,N,4]` (the uniforms of the reset that may follow step t, standing in for
    the reference's global-`random` draws, e.g. env_ball_balance_v0.py:193-194) finished envs auto-reset inside the step, as
    the reference's caller would, and `obs` holds the first observation of the new episode for those steps."""
    actions = np.asarray(actions, np.float32)
    T, N = actions.shape[:2]
    u = np.zeros((T, N), np.float32) if u is None else np.asarray(u, np.float32)
    tr = dict(format_version=FORMAT_VERSION, env=env_name, state0=_np(backend.get_state()).astype(np.float64), actions=actions, u=u)
    resets = np.zeros(N, int)      # auto-resets so far per env: `draws` may be a function (t, resets) -> [N,4], so that the
    used = []                      # k-th reset of an env consumes the k-th draw of a sequence, as a global RNG would
    cols = {k: [] for k in ('obs', 'reward', 'terminated', 'truncated', 'success', 'nticks', 'pairs', 'counts', 'state')}
    for t in range(T):
        d = None if draws is None else np.asarray(draws(t, resets) if callable(draws) else draws[t], np.float64)
        if d is not None:
            used.append(d)
        obs, rew, term, trunc, succ, nt = _step(backend, actions[t].astype(np.float64), u[t].astype(np.float64), d, draws is not None)
        resets += (np.asarray(term).astype(bool) | np.asarray(trunc).astype(bool)).astype(int)
        pairs, counts = (_np(x) for x in backend.contacts())
        wide = -np.ones((pairs.shape[0], PAIRS_WIDTH, 2), pairs.dtype)
        wide[:, :pairs.shape[1]] = pairs[:, :PAIRS_WIDTH]
        for k, v in zip(cols, (obs, rew, term, trunc, succ, nt, wide, np.minimum(counts, pairs.shape[1]), _np(backend.get_state()))):
            cols[k].append(np.array(v, copy=True))
    tr.update({k: np.stack(v) for k, v in cols.items()})
    if used:
        tr['draws'] = np.stack(used)
    return tr


def save(path, trace):
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in trace.items()})


def load(path):
    with np.load(path, allow_pickle=False) as z:
        tr = {k: z[k] for k in z.files}
    if int(tr['format_version']) != FORMAT_VERSION:
        raise ValueError('trace format %s, this reader understands %d' % (tr['format_version'], FORMAT_VERSION))
    tr['env'] = str(tr['env'])
    return tr


def replay(backend, trace, resync=True):
    """Re-run a trace on `backend` and return per-step differences.

    resync=True starts every step from the recorded state (per-step error, no chaotic growth); resync=False runs
    free from state0. Returns a dict of arrays of length T: max |obs| error, max |reward| error, fraction of envs with
    equal trip counts, equal flags, equal contact-pair sets, max |state| error."""
    T = trace['actions'].shape[0]
    out = {k: np.zeros(T) for k in ('obs_err', 'reward_err', 'same_ticks', 'same_flags', 'same_contacts', 'state_err')}
    backend.set_state(trace['state0'])
    draws = trace.get('draws')
    for t in range(T):
        if resync and t > 0:
            backend.set_state(trace['state'][t - 1])
        obs, rew, term, trunc, succ, nt = _step(backend, trace['actions'][t].astype(np.float64), trace['u'][t].astype(np.float64),
                                                None if draws is None else draws[t], draws is not None)
        pairs, counts = (_np(x) for x in backend.contacts())
        out['obs_err'][t] = np.abs(obs - trace['obs'][t]).max()
        out['reward_err'][t] = np.abs(rew - trace['reward'][t]).max()
        out['same_ticks'][t] = np.mean(np.all(np.asarray(nt) == trace['nticks'][t], axis=1))     # settle ticks of the action and ticks of the auto-reset
        out['same_flags'][t] = np.mean((np.asarray(term).astype(bool) == trace['terminated'][t].astype(bool)) &
                                       (np.asarray(succ).astype(bool) == trace['success'][t].astype(bool)))
        same = [set(map(tuple, pairs[i, :counts[i]].tolist())) == set(map(tuple, trace['pairs'][t][i, :trace['counts'][t][i]].tolist()))
                for i in range(len(counts))]
        out['same_contacts'][t] = np.mean(same)
        out['state_err'][t] = np.abs(_np(backend.get_state())[:, :77] - trace['state'][t][:, :77]).max()
    return out
