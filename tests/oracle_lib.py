"""ctypes binding of the CPU oracle (oracle/_build/liburgo.so). Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV_TYPES = {'BallBalance': 0, 'TwoBallsBalance': 1, 'CubesPush': 2, 'CubesGrasp': 3, 'Golf': 4}
MAXC = 16
STATE_SLOTS = 96
_lib = None


def build():
    subprocess.check_call(['make', '-s', '-C', os.path.join(ROOT, 'oracle')])


def lib():
    global _lib
    if _lib is None:
        so = os.path.join(ROOT, 'oracle', '_build', 'liburgo.so')
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.urgo_create.restype = C.c_void_p
        L.urgo_create.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_int]
        L.urgo_destroy.argtypes = [C.c_void_p]
        for f in ('urgo_reset', 'urgo_step', 'urgo_get_state', 'urgo_set_state', 'urgo_contacts', 'urgo_tick',
                  'urgo_last_contacts_detail', 'urgo_last_iters', 'urgo_reset_range', 'urgo_step_range'):
            getattr(L, f).restype = C.c_int
        L.urgo_total_ticks.restype = C.c_long
        L.urgo_total_ticks.argtypes = [C.c_void_p]
        L.urgo_geometric_distance_reward.restype = C.c_double
        L.urgo_geometric_distance_reward.argtypes = [C.c_double] * 3
        L.urgo_gripper_angle.restype = C.c_double
        L.urgo_gripper_angle.argtypes = [C.c_double]
        L.urgo_urand.restype = C.c_double
        L.urgo_urand.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]
        L.urgo_box_box.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _d(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


class OracleBatch:
    def __init__(self, env, num_envs, seed=0, offset=0):
        self.L = lib()
        self.t = ENV_TYPES[env] if isinstance(env, str) else int(env)
        self.n = num_envs
        self.h = C.c_void_p(self.L.urgo_create(self.t, num_envs, seed, offset))
        self.od = self.L.urgo_obs_dim(self.t)
        self.ad = self.L.urgo_act_dim(self.t)
        self.max_steps = self.L.urgo_max_steps(self.t)

    def __del__(self):
        if getattr(self, 'h', None):
            self.L.urgo_destroy(self.h)
            self.h = None

    def reset(self, mask=None, draws=None, lo=0, hi=None):
        obs = np.zeros((self.n, self.od))
        nt = np.zeros(self.n, dtype=np.int32)
        m = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
        dr = _d(draws)
        self.L.urgo_reset_range(self.h, lo, self.n if hi is None else hi, _p(m), _p(dr), _p(obs), _p(nt))
        return obs, nt

    def step(self, action, u=None, draws=None, autoreset=False, lo=0, hi=None):
        a = _d(action).reshape(self.n, self.ad)
        obs = np.zeros((self.n, self.od)); fin = np.zeros((self.n, self.od))
        rew = np.zeros(self.n)
        term = np.zeros(self.n, np.uint8); trunc = np.zeros(self.n, np.uint8); succ = np.zeros(self.n, np.uint8)
        nt = np.zeros((self.n, 2), np.int32)
        uu, dr = _d(u), _d(draws)
        self.L.urgo_step_range(self.h, lo, self.n if hi is None else hi, _p(a), _p(uu), _p(dr), _p(obs), _p(rew), _p(term), _p(trunc), _p(succ), _p(fin), _p(nt), int(autoreset))
        return dict(obs=obs, reward=rew, terminated=term.astype(bool), truncated=trunc.astype(bool),
                    success=succ.astype(bool), final_obs=fin, nticks=nt)

    def get_state(self):
        s = np.zeros((self.n, STATE_SLOTS)); self.L.urgo_get_state(self.h, _p(s)); return s

    def set_state(self, s):
        s = _d(s).reshape(self.n, STATE_SLOTS); self.L.urgo_set_state(self.h, _p(s))

    def contacts(self):
        pairs = np.zeros((self.n, MAXC, 2), np.int32); cnt = np.zeros(self.n, np.int32)
        self.L.urgo_contacts(self.h, _p(pairs), _p(cnt)); return pairs, cnt

    def tick(self, n=1):
        self.L.urgo_tick(self.h, n)

    def contacts_detail(self, env=0):
        out = np.zeros((MAXC, 11)); n = self.L.urgo_last_contacts_detail(self.h, env, _p(out)); return out[:n]

    def last_iters(self, env=0):
        return self.L.urgo_last_iters(self.h, env)

    def set_sparse(self, flag):
        self.L.urgo_set_sparse(self.h, int(flag))

    def total_ticks(self):
        return self.L.urgo_total_ticks(self.h)


def fk(q):
    out = np.zeros((19, 7)); q = _d(q); lib().urgo_fk(_p(q), _p(out)); return out


def ik(q, target7):
    out = np.zeros(6); q = _d(q); t = _d(target7); lib().urgo_ik(_p(q), _p(t), _p(out)); return out


def gripper_center(q):
    out = np.zeros(3); q = _d(q); lib().urgo_gripper_center(_p(q), _p(out)); return out


def forward_dynamics(q, qd, gz=-9.8):
    out = np.zeros(12); q = _d(q); qd = _d(qd); lib().urgo_forward_dynamics(_p(q), _p(qd), C.c_double(gz), _p(out)); return out


def minv(q):
    out = np.zeros((12, 12)); q = _d(q); lib().urgo_minv(_p(q), _p(out)); return out


def euler_from_quat(q):
    out = np.zeros(3); q = _d(q); lib().urgo_euler_from_quat(_p(q), _p(out)); return out


def quat_from_euler(e):
    out = np.zeros(4); e = _d(e); lib().urgo_quat_from_euler(_p(e), _p(out)); return out


def box_box(a, b, margin):
    out = np.zeros((4, 10)); a = _d(a); b = _d(b); n = lib().urgo_box_box(_p(a), _p(b), margin, _p(out)); return out[:n]
