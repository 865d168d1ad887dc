"""ctypes binding of the CPU oracle (oracle/_build/liburgo.so). Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV_TYPES = {'BallBalance': 0, 'TwoBallsBalance': 1, 'CubesPush': 2, 'CubesGrasp': 3, 'Golf': 4, 'CubesPushDiff': 5, 'GolfJoints': 6, 'Box': 7}
MAXC = 64          # width of the oracle's contact-pair export (urgo_export_maxc); the oracle itself keeps every point
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
                  'urgo_last_contacts_detail', 'urgo_last_iters', 'urgo_reset_range', 'urgo_step_range', 'urgo_prologue',
                  'urgo_epilogue', 'urgo_trace'):
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

    def step_parallel(self, action, u=None, draws=None, autoreset=False, threads=None):
        """step() with the envs spread over host threads (envs are independent; ctypes releases the GIL): same results."""
        from concurrent.futures import ThreadPoolExecutor
        threads = min(threads or (os.cpu_count() or 1), self.n)
        if threads <= 1:
            return self.step(action, u=u, draws=draws, autoreset=autoreset)
        cuts = np.linspace(0, self.n, threads + 1).astype(int)
        with ThreadPoolExecutor(threads) as ex:
            parts = list(ex.map(lambda i: self.step(action, u=u, draws=draws, autoreset=autoreset, lo=int(cuts[i]), hi=int(cuts[i + 1])), range(threads)))
        out = parts[0]
        for i in range(1, threads):
            sl = slice(int(cuts[i]), int(cuts[i + 1]))
            for k in out:
                out[k][sl] = parts[i][k][sl]
        return out

    def get_state(self):
        s = np.zeros((self.n, STATE_SLOTS)); self.L.urgo_get_state(self.h, _p(s)); return s

    def set_state(self, s):
        s = _d(s).reshape(self.n, STATE_SLOTS); self.L.urgo_set_state(self.h, _p(s))

    def contacts(self):
        pairs = np.zeros((self.n, MAXC, 2), np.int32); cnt = np.zeros(self.n, np.int32)
        assert self.L.urgo_export_maxc() == MAXC
        self.L.urgo_contacts(self.h, _p(pairs), _p(cnt)); return pairs, np.minimum(cnt, MAXC)

    def tick(self, n=1):
        self.L.urgo_tick(self.h, n)

    def contacts_detail(self, env=0):
        out = np.zeros((MAXC, 11)); n = self.L.urgo_last_contacts_detail(self.h, env, _p(out)); return out[:n]

    def settle_trace(self, env=0):
        """Summed |joint motion| of every tick of the env's last wait_until_stable."""
        out = np.zeros(480); n = self.L.urgo_settle_trace(self.h, env, _p(out), 480); return out[:n]

    def contact_margin(self, env=0):
        """Smallest |inequality slack| the narrowphase decided on in the env's last tick (how close the contact set came to another)."""
        self.L.urgo_contact_margin.restype = C.c_double
        return float(self.L.urgo_contact_margin(self.h, env))

    def last_iters(self, env=0):
        return self.L.urgo_last_iters(self.h, env)

    def set_sparse(self, flag):
        self.L.urgo_set_sparse(self.h, int(flag))

    def tick_hist(self):
        """Histograms over all ticks so far: solver rows, manifold points, joint rows, rolling-friction manifolds per tick."""
        r, c, j, t = np.zeros(192, np.int64), np.zeros(64, np.int64), np.zeros(32, np.int64), np.zeros(16, np.int64)
        self.L.urgo_tick_hist(self.h, _p(r), _p(c), _p(j), _p(t))
        return dict(rows=r, contacts=c, joint_rows=j, rolling=t)

    def total_ticks(self):
        return self.L.urgo_total_ticks(self.h)

    # ---- env logic without physics (tests/test_reference_env_logic.py)
    def prologue(self, action, u=0.0, env=0):
        a = _d(action); out = np.zeros(16)
        self.L.urgo_prologue(self.h, env, _p(a), C.c_double(u), _p(out))
        return dict(reward=out[0], targets=out[1:8].copy(), arm_armed=bool(out[8]), grip_armed=bool(out[9]))

    def epilogue(self, pairs, reward_so_far=0.0, env=0):
        pr = np.ascontiguousarray(np.asarray(pairs, dtype=np.int32).reshape(-1, 2))
        obs = np.zeros(self.od); rew = np.zeros(1)
        f = np.zeros(3, np.uint8)
        self.L.urgo_epilogue(self.h, env, C.c_double(reward_so_far), len(pr), _p(pr), _p(obs), _p(rew), _p(f[0:1]), _p(f[1:2]), _p(f[2:3]))
        return dict(obs=obs, reward=float(rew[0]), terminated=bool(f[0]), truncated=bool(f[1]), success=bool(f[2]))

    def set_trace(self, on=True):
        self.L.urgo_set_trace(self.h, int(on))

    def trace(self, env=0):
        out = np.zeros((256, 9)); n = self.L.urgo_trace(self.h, env, _p(out), 256); return out[:n]

    def kinematics(self, env=0):
        o = np.zeros(19); self.L.urgo_kinematics(self.h, env, _p(o))
        return dict(ee=o[0:7].copy(), link11=o[7:10].copy(), link16=o[10:13].copy(), gripper_center=o[13:16].copy(), club_foot=o[16:19].copy())

    def shape_table(self):
        """[(body, link)] per shape index: body 0 robot, k>0 free body k-1 (link -1 = its base), <0 static (link = which)."""
        out = []
        for i in range(self.L.urgo_num_shapes(self.t)):
            b, l = C.c_int(), C.c_int(); self.L.urgo_shape_info(self.t, i, C.byref(b), C.byref(l)); out.append((b.value, l.value))
        return out


def set_param(name, value):
    """Run-time copy of a recalled Bullet parameter (oracle/bullet_params.h: BulletParams); worlds created afterwards see it."""
    L = lib(); L.urgo_set_param.argtypes = [C.c_char_p, C.c_double]
    if L.urgo_set_param(name.encode(), float(value)) != 0:
        raise KeyError(name)


def reset_params():
    lib().urgo_reset_params()


def fk(q):
    out = np.zeros((19, 7)); q = _d(q); lib().urgo_fk(_p(q), _p(out)); return out


def ik(q, target7, base=None):
    out = np.zeros(6); q = _d(q); t = _d(target7)
    if base is None:
        lib().urgo_ik(_p(q), _p(t), _p(out))
    else:
        b = _d(np.asarray(base, dtype=np.float64)); lib().urgo_ik_base(_p(q), _p(t), _p(b), _p(out))
    return out


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


def settle_count(l1_motion_per_tick):
    a = _d(l1_motion_per_tick); L = lib(); L.urgo_settle_count.restype = C.c_int
    return L.urgo_settle_count(_p(a), len(a))


def model_facts():
    """Static facts of the oracle's robot / env model (see urgo_model_info in oracle/urgo_env.cpp)."""
    o = np.zeros(81); lib().urgo_model_info(_p(o))
    return dict(dof_link=o[0:12].astype(int).tolist(), effort=o[12:24].tolist(), velocity=o[24:36].tolist(), lower=o[36:48].tolist(),
                upper=o[48:60].tolist(), gear_child_link=o[60:65].astype(int).tolist(), gear_ratio=o[65:70].tolist(),
                rest_pose=o[70:76].tolist(), reset_arm_ticks=int(o[76]), settle_max_ticks=int(o[77]), settle_threshold=float(o[78]),
                club_zc=float(o[79]), ik_iters=int(o[80]))


def scene_facts(env):
    t = ENV_TYPES[env] if isinstance(env, str) else int(env)
    L = lib(); o = np.zeros(4); L.urgo_scene_info(t, _p(o))
    shapes = []
    for i in range(L.urgo_num_shapes(t)):
        b, l = C.c_int(), C.c_int(); L.urgo_shape_info(t, i, C.byref(b), C.byref(l))
        d = np.zeros(10); L.urgo_shape_detail(t, i, _p(d))
        shapes.append(dict(body=b.value, link=l.value, kind='box' if d[0] == 0 else 'sphere', half=d[1:4].tolist(), radius=float(d[1]),
                           centre=d[4:7].tolist(), mu=float(d[7]), restitution=float(d[8]), rolling=float(d[9])))
    return dict(gravity=float(o[0]), masses=o[1:4].tolist(), shapes=shapes)
