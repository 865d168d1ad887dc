"""Stable-Baselines3 `VecEnv` view of a device-resident batch.

The reference trains with SB3 on one pybullet env per process (`optuna_training.py:35-56`, `training.py`). SB3's
algorithms only need the `VecEnv` protocol, so the whole batch can be handed to them as one vectorised env:

    from urgym_b200.sb3_vec_env import URGymSB3VecEnv
    venv = URGymSB3VecEnv('URGym/CubesPush-v0', num_envs=4096, device=0, seed=0)
    model = SAC('MlpPolicy', venv, ...)          # stable_baselines3, when installed

SB3's conventions are followed: `reset()` -> float32 numpy observations `[N, obs_dim]`; `step_async(actions)` /
`step_wait()` -> `(obs, rewards, dones, infos)` with `dones = terminated | truncated`, finished envs already reset
(the observation returned is the first one of the next episode), and per-env `infos[i]` carrying
`terminal_observation`, `TimeLimit.truncated` and `is_success` where the reference provides it
(`env_cubes_push_v0.py:143`). When stable_baselines3 is importable the class derives from its `VecEnv`; otherwise it is
a duck-typed stand-in with the same methods. SB3 works on host numpy arrays, so this adapter costs one host round
trip per step (it uses the host-buffer entry `urgym_step_host`); `URGymVectorEnv` is the sync-free device API.
"""
import ctypes as C

import numpy as np

from . import _lib
from .envs import HAS_SUCCESS_INFO, SPECS, _spaces

try:
    from stable_baselines3.common.vec_env import VecEnv as _Base
except Exception:  # stable_baselines3 is not installed in this image
    _Base = object


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class URGymSB3VecEnv(_Base):
    def __init__(self, env_id, num_envs, device=0, seed=0, reward_type='dense', env_index_offset=0):
        if env_id not in SPECS:
            raise ValueError('%s is not backed by urgym_b200' % env_id)
        self.env_id, self.name = env_id, SPECS[env_id][0]
        self.L = _lib.lib()
        t = _lib.ENV_TYPES[self.name]
        h = C.c_void_p()
        _lib.check(self.L.urgym_create(t, int(num_envs), int(device), int(seed), int(env_index_offset), int(reward_type == 'sparse'), C.byref(h)))
        self.h = h
        obs_space, act_space = _spaces(env_id)
        if _Base is object:
            self.num_envs, self.observation_space, self.action_space = int(num_envs), obs_space, act_space
        else:
            super().__init__(int(num_envs), obs_space, act_space)
        n, od = self.num_envs, obs_space.shape[0]
        self._obs = np.zeros((n, od), np.float32); self._final = np.zeros((n, od), np.float32)
        self._rew = np.zeros(n, np.float32)
        self._term, self._trunc, self._succ = (np.zeros(n, np.uint8) for _ in range(3))
        self._nticks = np.zeros((n, 2), np.int32)
        self._actions = None
        self.render_mode = None

    # ---- VecEnv protocol
    def reset(self):
        _lib.check(self.L.urgym_reset_host(self.h, None, None, _p(self._obs), None))
        return self._obs.copy()

    def step_async(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.float32).reshape(self.num_envs, self.action_space.shape[0])
        self._actions = np.clip(a, self.action_space.low, self.action_space.high)   # SB3 clips to the Box as well

    def step_wait(self):
        _lib.check(self.L.urgym_step_host(self.h, _p(self._actions), None, None, _p(self._obs), _p(self._rew), _p(self._term), _p(self._trunc),
                                          _p(self._succ), _p(self._final), _p(self._nticks), 1))
        dones = (self._term | self._trunc).astype(bool)
        infos = [{} for _ in range(self.num_envs)]
        with_success = self.name in HAS_SUCCESS_INFO
        for i in np.nonzero(dones)[0]:
            infos[i]['terminal_observation'] = self._final[i].copy()
            infos[i]['TimeLimit.truncated'] = bool(self._trunc[i] and not self._term[i])
        if with_success:
            for i in range(self.num_envs):
                infos[i]['is_success'] = bool(self._succ[i])
        return self._obs.copy(), self._rew.copy(), dones, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self):
        if getattr(self, 'h', None):
            self.L.urgym_destroy(self.h)
            self.h = None

    def seed(self, seed=None):
        return [None] * self.num_envs          # the batch is seeded at construction (counter-based RNG per env index)

    def get_attr(self, attr_name, indices=None):
        return [getattr(self, attr_name)] * len(self._indices(indices))

    def set_attr(self, attr_name, value, indices=None):
        setattr(self, attr_name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        raise NotImplementedError('per-env methods do not exist on a device batch (%s)' % method_name)

    def env_is_wrapped(self, wrapper_class, indices=None):
        return [False] * len(self._indices(indices))

    def get_images(self):
        raise NotImplementedError('no rendering: unavailable on this box')

    def _indices(self, indices):
        if indices is None:
            return list(range(self.num_envs))
        return [indices] if isinstance(indices, int) else list(indices)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
