"""Device-resident batch of URGym worlds: thin torch-facing wrapper over the C ABI (no host sync on the step path)."""
import ctypes as C

import torch

from . import _lib
from ._lib import ENV_TYPES, MAX_CONTACTS, NUM_LINKS, PAIR_WORDS, STATE_FLOATS, check


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


_DTYPE = {'f32': torch.float32, 'f64': torch.float64}


def _real(t, dev, shape=None, dtype=torch.float32):
    if t is None:
        return None
    t = torch.as_tensor(t, dtype=dtype, device=dev).contiguous()
    if shape is not None:
        t = t.reshape(shape)
    return t


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


class EnvBatch:
    """`num_envs` independent instances of one env type on one GPU.

    precision='f32' is the product (liburgym_b200.so); 'f64' loads the parity build of the same kernels
    (liburgym_b200_f64.so), whose buffers are float64 (tests only: integer outputs against the fp64 oracle)."""

    def __init__(self, env, num_envs, device=0, seed=0, env_index_offset=0, reward_sparse=False, precision='f32'):
        if not torch.cuda.is_available():
            raise _lib.UrgymError('CUDA is not available: urgym_b200 has no CPU path')
        self.L = _lib.lib(precision)
        self.dtype = _DTYPE[precision]
        self.env_type = ENV_TYPES[env] if isinstance(env, str) else int(env)
        self.num_envs = int(num_envs)
        self.device = torch.device('cuda', device if isinstance(device, int) else torch.device(device).index or 0)
        self.obs_dim = self.L.urgym_obs_dim(self.env_type)
        self.act_dim = self.L.urgym_act_dim(self.env_type)
        self.max_episode_steps = self.L.urgym_max_episode_steps(self.env_type)
        h = C.c_void_p()
        check(self.L.urgym_create(self.env_type, self.num_envs, self.device.index, seed, env_index_offset, int(reward_sparse), C.byref(h)), self.L)
        self.h = h
        n, d = self.num_envs, self.device
        self.obs = torch.zeros(n, self.obs_dim, device=d, dtype=self.dtype)
        self.final_obs = torch.zeros(n, self.obs_dim, device=d, dtype=self.dtype)
        self.reward = torch.zeros(n, device=d, dtype=self.dtype)
        self.terminated = torch.zeros(n, dtype=torch.uint8, device=d)
        self.truncated = torch.zeros(n, dtype=torch.uint8, device=d)
        self.success = torch.zeros(n, dtype=torch.uint8, device=d)
        self.nticks = torch.zeros(n, 2, dtype=torch.int32, device=d)
        self.reset_ticks = torch.zeros(n, dtype=torch.int32, device=d)

    def close(self):
        if getattr(self, 'h', None) is not None and self.h:
            self.L.urgym_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self, mask=None, draws=None):
        d = self.device
        m = None if mask is None else torch.as_tensor(mask, device=d).to(torch.uint8).contiguous()
        dr = _real(draws, d, (self.num_envs, 4), self.dtype)
        check(self.L.urgym_reset(self.h, _ptr(m), _ptr(dr), _ptr(self.obs), _ptr(self.reset_ticks), _stream(d)), self.L)
        return self.obs

    def step(self, action, u=None, draws=None, autoreset=True):
        d = self.device
        a = _real(action, d, (self.num_envs, self.act_dim), self.dtype)
        uu = _real(u, d, (self.num_envs,), self.dtype)
        dr = _real(draws, d, (self.num_envs, 4), self.dtype)
        check(self.L.urgym_step(self.h, _ptr(a), _ptr(uu), _ptr(dr), _ptr(self.obs), _ptr(self.reward), _ptr(self.terminated),
                                _ptr(self.truncated), _ptr(self.success), _ptr(self.final_obs), _ptr(self.nticks), int(autoreset), _stream(d)), self.L)
        return self.obs, self.reward, self.terminated, self.truncated, self.success

    def step_logic(self, action, pairs, counts, u=None):
        """step() with the physics skipped and the contact set given (include/urgym_b200.h: urgym_step_logic)."""
        d = self.device
        a = _real(action, d, (self.num_envs, self.act_dim), self.dtype)
        uu = _real(u, d, (self.num_envs,), self.dtype)
        pr = torch.as_tensor(pairs, dtype=torch.int32, device=d).reshape(self.num_envs, MAX_CONTACTS, 2).contiguous()
        cn = torch.as_tensor(counts, dtype=torch.int32, device=d).reshape(self.num_envs).contiguous()
        check(self.L.urgym_step_logic(self.h, _ptr(a), _ptr(uu), _ptr(pr), _ptr(cn), _ptr(self.obs), _ptr(self.reward), _ptr(self.terminated),
                                      _ptr(self.truncated), _ptr(self.success), _stream(d)), self.L)
        return self.obs, self.reward, self.terminated, self.truncated, self.success

    def tick(self, n=1):
        check(self.L.urgym_tick(self.h, int(n), _stream(self.device)), self.L)

    def get_state(self):
        s = torch.empty(self.num_envs, STATE_FLOATS, device=self.device, dtype=self.dtype)
        check(self.L.urgym_get_state(self.h, _ptr(s), _stream(self.device)), self.L)
        return s

    def set_state(self, s):
        s = _real(s, self.device, (self.num_envs, STATE_FLOATS), self.dtype)
        check(self.L.urgym_set_state(self.h, _ptr(s), _stream(self.device)), self.L)
        torch.cuda.current_stream(self.device).synchronize()

    def contacts(self):
        pairs = torch.empty(self.num_envs, MAX_CONTACTS, 2, dtype=torch.int32, device=self.device)
        cnt = torch.empty(self.num_envs, dtype=torch.int32, device=self.device)
        check(self.L.urgym_contacts(self.h, _ptr(pairs), _ptr(cnt), _stream(self.device)), self.L)
        return pairs, cnt

    def contact_pairs(self):
        """Set of (shape_a, shape_b) in contact per env after the last step (urgym_contact_pairs: independent of the point capacity)."""
        mask = torch.empty(self.num_envs, PAIR_WORDS, dtype=torch.int32, device=self.device)
        check(self.L.urgym_contact_pairs(self.h, _ptr(mask), _stream(self.device)), self.L)
        m = mask.cpu().numpy().view('uint32')
        if not hasattr(self, '_pairs'):
            self._pairs = _lib.pair_table(self.env_type, self.L.precision)
        return [{self._pairs[p] for p in range(len(self._pairs)) if (int(m[i, p >> 5]) >> (p & 31)) & 1} for i in range(self.num_envs)]

    def stats(self):
        out = (C.c_double * 8)()
        check(self.L.urgym_stats(self.h, out), self.L)
        return dict(zip(('episodes', 'return_sum', 'length_sum', 'successes', 'env_steps', 'ticks', 'capacity_overflow_ticks'), list(out)[:7]))


def forward_kinematics(q, device=0, precision='f32'):
    """q [n,12] -> link poses [n,19,7] (xyz, quat xyzw); replaces getLinkState for every link."""
    dev = torch.device('cuda', device)
    q = _real(q, dev, dtype=_DTYPE[precision]).reshape(-1, 12)
    out = torch.empty(q.shape[0], NUM_LINKS, 7, device=dev, dtype=q.dtype)
    L = _lib.lib(precision)
    check(L.urgym_fk(device, q.shape[0], _ptr(q), _ptr(out), _stream(dev)), L)
    return out


def inverse_kinematics(q, target, device=0, precision='f32'):
    """q [n,12], target [n,7] -> arm joint targets [n,6]; replaces calculateInverseKinematics (robot.py:119-121)."""
    dev = torch.device('cuda', device)
    q = _real(q, dev, dtype=_DTYPE[precision]).reshape(-1, 12)
    t = _real(target, dev, dtype=q.dtype).reshape(-1, 7)
    out = torch.empty(q.shape[0], 6, device=dev, dtype=q.dtype)
    L = _lib.lib(precision)
    check(L.urgym_ik(device, q.shape[0], _ptr(q), _ptr(t), _ptr(out), _stream(dev)), L)
    return out


def gripper_center(q, device=0, precision='f32'):
    dev = torch.device('cuda', device)
    q = _real(q, dev, dtype=_DTYPE[precision]).reshape(-1, 12)
    out = torch.empty(q.shape[0], 3, device=dev, dtype=q.dtype)
    L = _lib.lib(precision)
    check(L.urgym_gripper_center(device, q.shape[0], _ptr(q), _ptr(out), _stream(dev)), L)
    return out


def step_static(env, state_in, action, out, u=None, draws=None, seed=0, env_index_offset=0, reward_sparse=False, autoreset=True,
                precision='f32'):
    """One gym step through the allocation-free entry (include/urgym_b200.h: urgym_step_static). Every buffer is the caller's:
    `state_in` [n,96] and the dict `out` of preallocated device tensors {'state','obs','reward','terminated','truncated',
    'success'} plus optional 'final_obs', 'nticks' [n,2] int32 and 'stats' [8] float64. No tensor is created here when the
    inputs already are contiguous device tensors of the library's dtype."""
    L = _lib.lib(precision)
    env_type = ENV_TYPES[env] if isinstance(env, str) else int(env)
    dev = state_in.device
    n = state_in.shape[0]
    dt = _DTYPE[precision]
    a = _real(action, dev, (n, -1), dt)
    check(L.urgym_step_static(env_type, dev.index or 0, n, seed, env_index_offset, int(reward_sparse), _ptr(state_in), _ptr(a),
                              _ptr(_real(u, dev, (n,), dt)), _ptr(_real(draws, dev, (n, 4), dt)), _ptr(out['state']), _ptr(out['obs']),
                              _ptr(out['reward']), _ptr(out['terminated']), _ptr(out['truncated']), _ptr(out['success']),
                              _ptr(out.get('final_obs')), _ptr(out.get('nticks')), _ptr(out.get('stats')), int(autoreset), _stream(dev)), L)
    return out
