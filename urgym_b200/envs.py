"""Host-side mirror of the reference's Gymnasium interface for every env it registers (the five of the
north star, the two variants CubesPushDiff and GolfJoints, and Box).

Reference: /root/reference/urgym/envs/__init__.py:1-50 (ids, max_episode_steps) and the env classes, e.g.
/root/reference/urgym/envs/env_ball_balance_v0.py:34-173 (`__init__(camera=None, reward_type='dense',
render_mode='human')`, `reset(*, seed, options) -> (obs float64[n], {})`, `step(action) -> (obs, reward, terminated,
truncated, info)`, `close()`, `observation_space`, `action_space`).

Two entry points:
  * `make(env_id, **kwargs)` / the classes `BallBalance`, ... : a batch of ONE env with numpy in / out, float64
    observations as the declared spaces say, and the TimeLimit wrapper's `truncated` applied (what `gym.make`
    returns for the reference). When `gymnasium` is importable the same ids are registered there too.
  * `URGymVectorEnv(env_id, num_envs, device, seed)`: device-resident batch, torch CUDA tensors in / out, in-kernel
    auto-reset (same-step: `obs` of a finished env is the first observation of its next episode,
    `info['final_observation']` holds the last one of the finished episode), no host synchronisation on the step path.

Everything physical happens in liburgym_b200.so through the C ABI (include/urgym_b200.h). There is no CPU fallback.
"""
import math
import os

import numpy as np

from ._lib import ENV_TYPES

try:  # the reference depends on gymnasium==0.29.1 (setup.py:9); it is optional here
    import gymnasium as _gym
    from gymnasium.spaces import Box as BoxSpace
except Exception:  # minimal stand-in used only when gymnasium cannot be imported
    _gym = None

    class BoxSpace:
        def __init__(self, low, high, dtype=np.float32):
            self.low = np.asarray(low, dtype=dtype); self.high = np.asarray(high, dtype=dtype)
            self.dtype = np.dtype(dtype); self.shape = self.low.shape
            self._rng = np.random.default_rng()

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)

        def sample(self):
            return self._rng.uniform(self.low, self.high).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

        def __repr__(self):
            return 'Box(%s, %s, %s, %s)' % (self.low.min(), self.high.max(), self.shape, self.dtype)

PI = math.pi
# id -> (class name, max_episode_steps, obs low/high, act low/high); urgym/envs/__init__.py:10-44 and the constructors
SPECS = {
    'URGym/BallBalance-v0': ('BallBalance', 500, ([-PI] * 3 + [-1.0] * 3, [PI] * 3 + [1.0] * 3), ([-0.1] * 3, [0.1] * 3)),
    'URGym/TwoBallsBalance-v0': ('TwoBallsBalance', 500, ([-PI] * 3 + [-1.0] * 6, [PI] * 3 + [1.0] * 6), ([-0.1] * 3, [0.1] * 3)),
    'URGym/CubesPush-v0': ('CubesPush', 50, ([-1.0] * 9, [1.0] * 9), ([-0.1] * 3, [0.1] * 3)),
    'URGym/CubesGrasp-v0': ('CubesGrasp', 20, ([0] * 2 + [-1.0] * 3 + [-1] * 4 + [0] + [-1.0] * 3 + [-1.0] * 4, [1] * 2 + [1.0] * 3 + [1.0] * 4 + [1] + [1.0] * 3 + [1.0] * 4),
                            ([0] * 2 + [-PI] * 6 + [0], [1] * 2 + [PI] * 6 + [1])),
    'URGym/Golf-v0': ('Golf', 100, ([-1.0] * 9, [1.0] * 9), ([-0.1] * 3, [0.1] * 3)),
    # variants: same world, other observation / action mapping (env_cubes_push_diff_v0.py:209-211, env_golf_joints_v0.py:60-63)
    'URGym/CubesPushDiff-v0': ('CubesPushDiff', 50, ([-1.0] * 9, [1.0] * 9), ([-0.1] * 3, [0.1] * 3)),
    'URGym/GolfJoints-v0': ('GolfJoints', 100, ([-2 * PI] * 6 + [-1.0] * 9, [2 * PI] * 6 + [1.0] * 9), ([-0.1] * 6, [0.1] * 6)),
    # env_box_v0.py:40-44 (observation: end-effector position + quaternion; action: pose increment + gripper opening length);
    # max_episode_steps 200 (urgym/envs/__init__.py:3-8)
    'URGym/Box-v0': ('Box', 200, ([-1.0] * 3 + [-PI] * 4, [1.0] * 3 + [PI] * 4), ([-0.05] * 3 + [-PI / 10] * 4 + [0.0], [0.05] * 3 + [PI / 10] * 4 + [0.085])),
}
UNSUPPORTED = ()   # every id the reference registers is backed
HAS_SUCCESS_INFO = ('CubesPush', 'CubesGrasp', 'Golf', 'CubesPushDiff', 'GolfJoints', 'Box')  # info={'is_success': ...} (env_cubes_push_v0.py:143 etc.)


def _spaces(env_id):
    _, _, (ol, oh), (al, ah) = SPECS[env_id]
    return (BoxSpace(low=np.array(ol, dtype=np.float64), high=np.array(oh, dtype=np.float64), dtype=np.float64),
            BoxSpace(low=np.array(al, dtype=np.float32), high=np.array(ah, dtype=np.float32), dtype=np.float32))


class URGymVectorEnv:
    """`num_envs` independent instances of one env on one GPU; tensors stay on the device."""

    def __init__(self, env_id, num_envs, device=0, seed=None, reward_type='dense', env_index_offset=0, render_mode=None, camera=None):
        from .batch import EnvBatch  # imports torch; needs CUDA
        if env_id not in SPECS:
            raise ValueError('%s is not backed by urgym_b200 (supported: %s)' % (env_id, ', '.join(SPECS)))
        self.env_id, self.name, self.max_episode_steps = env_id, SPECS[env_id][0], SPECS[env_id][1]
        self.num_envs = int(num_envs)
        self.single_observation_space, self.single_action_space = _spaces(env_id)
        self.batch = EnvBatch(self.name, self.num_envs, device=device, seed=_fresh_seed() if seed is None else seed, env_index_offset=env_index_offset,
                              reward_sparse=(reward_type == 'sparse'))
        self.device = self.batch.device

    def reset(self, *, seed=None, options=None):
        """All envs (or options['mask']) restart. options['draws']: float[N,4] uniforms replacing the RNG draws."""
        options = options or {}
        obs = self.batch.reset(mask=options.get('mask'), draws=options.get('draws'))
        return obs, {'reset_ticks': self.batch.reset_ticks}

    def step(self, actions, u=None, draws=None):
        b = self.batch
        obs, rew, term, trunc, succ = b.step(actions, u=u, draws=draws, autoreset=True)
        info = {'final_observation': b.final_obs, 'nticks': b.nticks}
        if self.name in HAS_SUCCESS_INFO:
            info['is_success'] = succ.bool()
        return obs, rew, term.bool(), trunc.bool(), info

    def episode_statistics(self, all_reduce=False):
        """Episode counters accumulated on the device; with all_reduce=True they are summed over all ranks (NCCL)."""
        from .shard import reduce_stats
        st = self.batch.stats()
        return reduce_stats(st, device=self.device) if all_reduce else st

    def close(self):
        self.batch.close()


def _fresh_seed():
    """The reference draws its spawn positions from the unseeded global `random` (env_ball_balance_v0.py:193-194): two envs
    built without a seed must not replay the same episodes."""
    return int.from_bytes(os.urandom(7), 'little')


_EnvBase = _gym.Env if _gym is not None else object   # gym.make() needs a gymnasium.Env (spec, unwrapped, the passive checker)


class _SingleEnv(_EnvBase):
    """Batch-of-one env with the reference's single-env signature (numpy in / out, float64 observations)."""
    env_id = None
    metadata = {'render_modes': ['human', 'train']}

    def __init__(self, camera=None, reward_type='dense', render_mode='human', device=0, seed=None):
        # render_mode='human' opens a pybullet GUI in the reference (env_ball_balance_v0.py:45-57); there is no GUI
        # here, the argument is accepted and ignored. `camera` is always None in the reference.
        import torch
        from .batch import EnvBatch
        self._torch = torch
        self.name, self.max_episode_steps = SPECS[self.env_id][0], SPECS[self.env_id][1]
        self.observation_space, self.action_space = _spaces(self.env_id)
        self.reward_type = reward_type
        self.render_mode = render_mode
        self.batch = EnvBatch(self.name, 1, device=device, seed=_fresh_seed() if seed is None else seed, reward_sparse=(reward_type == 'sparse'))

    def reset(self, *, seed=None, options=None):
        # as in the reference, `seed` reaches only Env.np_random (super().reset), which the envs never draw from
        if _gym is not None:
            super().reset(seed=seed)
        draws = (options or {}).get('draws')
        obs = self.batch.reset(draws=None if draws is None else np.asarray(draws, dtype=np.float32).reshape(1, 4))
        return obs[0].double().cpu().numpy(), {}

    def step(self, action, u=None):
        a = np.asarray(action, dtype=np.float32).reshape(1, -1)
        uu = None if u is None else np.asarray([u], dtype=np.float32)
        obs, rew, term, trunc, succ = self.batch.step(a, u=uu, autoreset=False)
        info = {'is_success': bool(succ[0].item())} if self.name in HAS_SUCCESS_INFO else {}
        return obs[0].double().cpu().numpy(), float(rew[0].item()), bool(term[0].item()), bool(trunc[0].item()), info

    def close(self):
        self.batch.close()


def _mk(env_id):
    return type(SPECS[env_id][0], (_SingleEnv,), {'env_id': env_id, '__doc__': 'urgym_b200 backend of %s' % env_id})


BallBalance = _mk('URGym/BallBalance-v0')
TwoBallsBalance = _mk('URGym/TwoBallsBalance-v0')
CubesPush = _mk('URGym/CubesPush-v0')
CubesGrasp = _mk('URGym/CubesGrasp-v0')
Golf = _mk('URGym/Golf-v0')
CubesPushDiff = _mk('URGym/CubesPushDiff-v0')
GolfJoints = _mk('URGym/GolfJoints-v0')


class Box(_SingleEnv):
    """urgym_b200 backend of URGym/Box-v0 (env_box_v0.py:21: `BoxManipulation(button_touch_mode='any', camera=None, render_mode='human')`).
    Only the registered default button_touch_mode='any' is backed (the button counts as pressed whoever pushed it down)."""
    env_id = 'URGym/Box-v0'

    def __init__(self, button_touch_mode='any', camera=None, render_mode='human', device=0, seed=None):
        if button_touch_mode != 'any':
            raise NotImplementedError("button_touch_mode=%r: only 'any' (the registered default, urgym/envs/__init__.py:7) is backed" % (button_touch_mode,))
        super().__init__(camera=camera, render_mode=render_mode, device=device, seed=seed)


BoxManipulation = Box
_CLASSES = {c.env_id: c for c in (BallBalance, TwoBallsBalance, CubesPush, CubesGrasp, Golf, CubesPushDiff, GolfJoints, Box)}


def make(env_id, **kwargs):
    """Stand-in for gym.make('URGym/...-v0') (the TimeLimit wrapper's truncation is applied by the backend)."""
    if env_id in UNSUPPORTED:
        raise NotImplementedError('%s is registered by the reference but out of scope for this backend (SURVEY.md §2)' % env_id)
    if env_id not in _CLASSES:
        raise KeyError(env_id)
    return _CLASSES[env_id](**kwargs)


def register_all():
    """Register the ids with gymnasium when it is installed (same ids as urgym/envs/__init__.py; the TimeLimit is applied by the
    backend, so no max_episode_steps is passed). gymnasium is not installed in this image or on the GPU box: this path is
    untested here (INTEGRATION.md says so)."""
    if _gym is None:
        return False
    from gymnasium.envs.registration import register, registry
    for env_id, (cls, _, _, _) in SPECS.items():
        if env_id not in registry:
            register(id=env_id, entry_point='urgym_b200.envs:%s' % cls)  # truncation is applied by the backend
    return True


assert set(SPECS[k][0] for k in SPECS) == set(ENV_TYPES)
