#!/usr/bin/env python
"""Generate tests/golden/reference_env_logic.json by RUNNING the reference's own Python (read-only, /root/reference).

pybullet, gymnasium and cv2 are not installable in this image, so the reference cannot simulate here. Its env classes
are nevertheless plain Python around `pybullet as p`: this script imports them UNMODIFIED with `sys.modules` stand-ins
for those three packages, where the `pybullet` stand-in is a *scripted world* (ScriptedBullet below) that answers the
queries the env code makes (joint states, link states, AABBs, base poses, contact points) from a table the script
fills, and logs every command the env code issues (resetJointState, setJointMotorControl2, calculateInverseKinematics,
createMultiBody, loadURDF, changeDynamics, setGravity, createConstraint, stepSimulation, ...).

What this pins (everything except rigid-body dynamics, which is Bullet's C++ and absent here):
  * registry: ids, entry points, max_episode_steps                         (urgym/envs/__init__.py)
  * observation / action spaces of the five envs                          (env constructors)
  * helper math: geometric_distance_reward, quaternion_multiply, normalize_quaternion, rotate_quaternion
                                                                          (urgym/base/utilities.py)
  * robot set-up: controllable joints, rest pose, gear constraints, gripper length->angle, motor limits
                                                                          (urgym/base/robot.py)
  * scene construction: gravity, bodies, shapes, masses, friction changes  (env constructors, create_* helpers)
  * reset(): the exact command sequence and spawn offsets for seeded draws (env reset())
  * wait_until_stable(): trip counts for scripted joint-motion decays      (env wait_until_stable())
  * step(): for scripted world states — the commanded IK target / joint targets, and the returned observation,
    reward, terminated, info                                              (env step() and its predicates)

The scripted world states come from the CPU oracle (so the kinematic answers are self-consistent); the *outputs*
recorded are the reference's. tests/test_reference_env_logic.py replays every case through the oracle's
prologue / epilogue entry points and compares.

Run here (needs /root/reference; the GPU box only sees the committed JSON):
    python tests/golden/make_golden.py
"""
import contextlib
import io
import json
import math
import os
import random
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get('URGYM_REFERENCE', '/root/reference')
sys.path.insert(0, os.path.join(ROOT, 'tests'))
sys.path.insert(0, os.path.join(ROOT, 'tools'))


# ----------------------------------------------------------------------------------------------- stand-in packages
def _jsonable(x):
    if isinstance(x, (np.floating, float)):
        return float(x)
    if isinstance(x, (np.integer, int)) and not isinstance(x, bool):
        return int(x)
    if isinstance(x, (bool, str)) or x is None:
        return x
    if isinstance(x, bytes):
        return x.decode()
    if isinstance(x, np.ndarray):
        return [_jsonable(v) for v in x.tolist()]
    if isinstance(x, (list, tuple)):
        return [_jsonable(v) for v in x]
    if isinstance(x, dict):
        return {str(k): _jsonable(v) for k, v in x.items()}
    return repr(x)


def quat_from_euler(e):  # roll, pitch, yaw -> xyzw (ZYX convention, what getQuaternionFromEuler documents)
    r, p, y = [float(v) for v in e]
    cr, sr, cp, sp, cy, sy = math.cos(r / 2), math.sin(r / 2), math.cos(p / 2), math.sin(p / 2), math.cos(y / 2), math.sin(y / 2)
    return (sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy)


def euler_from_quat(q):  # xyzw -> roll, pitch, yaw
    x, y, z, w = [float(v) for v in q]
    sarg = -2 * (x * z - w * y)
    if sarg <= -0.99999:
        return (0.0, -0.5 * math.pi, 2 * math.atan2(x, -y))
    if sarg >= 0.99999:
        return (0.0, 0.5 * math.pi, 2 * math.atan2(-x, y))
    return (math.atan2(2 * (y * z + w * x), w * w - x * x - y * y + z * z), math.asin(sarg),
            math.atan2(2 * (x * y + w * z), w * w + x * x - y * y - z * z))


class ScriptedBullet(types.ModuleType):
    """Scripted world with pybullet's call signatures. Queries answer from tables; commands are logged."""
    GUI, DIRECT = 1, 2
    JOINT_REVOLUTE, JOINT_PRISMATIC, JOINT_SPHERICAL, JOINT_PLANAR, JOINT_FIXED, JOINT_GEAR = 0, 1, 2, 3, 4, 6
    POSITION_CONTROL, VELOCITY_CONTROL, TORQUE_CONTROL = 2, 0, 1
    GEOM_SPHERE, GEOM_BOX, GEOM_CYLINDER, GEOM_MESH = 2, 3, 4, 5
    COV_ENABLE_GUI, URDF_ENABLE_CACHED_GRAPHICS_SHAPES = 1, 1024
    URDF_USE_SELF_COLLISION, URDF_MERGE_FIXED_LINKS = 8, 512   # pybullet's values (env_box_v0.py:85)

    def __init__(self, robot_joints):
        super().__init__('pybullet')
        self.robot_joints = robot_joints
        self.clear()

    def clear(self):
        self.log = []
        self.bodies = {}       # id -> dict(kind=..., pos=(..), orn=(..))
        self.next_id = 0
        self.shapes = []
        self.joint_pos = {}    # (body, joint) -> position
        self.link_state = {}   # (body, link) -> (comPos, comOrn, locPos, locOrn, framePos, frameOrn)
        self.aabb = {}         # (body, link) -> (min3, max3)
        self.contacts = []     # (bodyA, linkA, bodyB, linkB)
        self.ik_result = tuple([0.0] * 12)
        self.nsteps = 0
        self.on_step = None

    def _rec(self, name, *a, **k):
        self.log.append([name, _jsonable(a), _jsonable(k)])

    def _new_body(self, kind, pos, orn, **extra):
        i = self.next_id
        self.next_id += 1
        self.bodies[i] = dict(kind=kind, pos=tuple(float(v) for v in pos), orn=tuple(float(v) for v in orn), **extra)
        return i

    # ---- commands (logged)
    def connect(self, *a, **k): self._rec('connect', *a, **k); return 0
    def disconnect(self, *a, **k): self._rec('disconnect')
    def setAdditionalSearchPath(self, *a, **k): pass
    def setGravity(self, *a, **k): self._rec('setGravity', *a, **k)
    def resetDebugVisualizerCamera(self, *a, **k): pass
    def configureDebugVisualizer(self, *a, **k): pass
    def addUserDebugParameter(self, *a, **k): return 0
    def readUserDebugParameter(self, *a, **k): return 0.0
    def loadTexture(self, *a, **k): return 0
    def changeVisualShape(self, *a, **k): pass
    def createVisualShape(self, *a, **k): return -1

    def loadURDF(self, fileName, basePosition=(0, 0, 0), baseOrientation=(0, 0, 0, 1), **k):
        name = os.path.basename(fileName)
        i = self._new_body(name, basePosition, baseOrientation)
        self._rec('loadURDF', name, basePosition, baseOrientation, body=i, **k)
        return i

    def createCollisionShape(self, shapeType=None, **k):
        self.shapes.append(dict(shapeType=shapeType, **_jsonable(k)))
        self._rec('createCollisionShape', shapeType, shape=len(self.shapes) - 1, **k)
        return len(self.shapes) - 1

    def createMultiBody(self, baseMass=0, baseCollisionShapeIndex=-1, baseVisualShapeIndex=-1, basePosition=(0, 0, 0),
                        baseOrientation=(0, 0, 0, 1), **k):
        i = self._new_body('multibody', basePosition, baseOrientation)
        if baseCollisionShapeIndex >= 0 or baseMass:  # bodies without collision are markers
            self._rec('createMultiBody', baseMass=baseMass, shape=baseCollisionShapeIndex, basePosition=basePosition,
                      baseOrientation=baseOrientation, body=i,
                      **{kk: v for kk, v in k.items() if 'Visual' not in kk})
        return i

    def removeBody(self, i): self._rec('removeBody', i); self.bodies.pop(i, None)
    def changeDynamics(self, *a, **k): self._rec('changeDynamics', *a, **k)
    def createConstraint(self, *a, **k): self._rec('createConstraint', *a, **k); return len(self.log)
    def changeConstraint(self, *a, **k): self._rec('changeConstraint', *a[1:], **k)
    def resetJointState(self, body, joint, value=None, *a, targetValue=None, **k):
        value = targetValue if value is None else value   # env_box_v0.py:173-174 passes it by keyword
        self._rec('resetJointState', joint, value); self.joint_pos[(body, joint)] = float(value)

    def setJointMotorControl2(self, body, joint, mode, *a, **k):
        if a:
            k = dict(k, targetPosition=a[0])
        self._rec('setJointMotorControl2', joint, mode, **k)

    def stepSimulation(self):
        self.nsteps += 1
        self._rec('stepSimulation')
        if self.on_step:
            self.on_step(self)

    def calculateInverseKinematics(self, body, link, pos, orn=None, *a, **k):
        self._rec('calculateInverseKinematics', link, pos, orn, nullspace=[list(x) for x in a], **k)
        return self.ik_result

    # ---- queries (answered from the tables)
    def getNumJoints(self, body): return len(self.robot_joints)
    def getJointInfo(self, body, i): return self.robot_joints[i]
    def getJointState(self, body, joint): return (self.joint_pos.get((body, joint), 0.0), 0.0, (0.0,) * 6, 0.0)
    def getLinkState(self, body, link, *a, **k):
        z, i = (0.0, 0.0, 0.0), (0.0, 0.0, 0.0, 1.0)
        return self.link_state.get((body, link), (z, i, z, i, z, i))

    def getAABB(self, body, link=-1): return self.aabb[(body, link)]
    def getBasePositionAndOrientation(self, body): b = self.bodies[body]; return b['pos'], b['orn']
    def getQuaternionFromEuler(self, e): return quat_from_euler(e)
    def getEulerFromQuaternion(self, q): return euler_from_quat(q)

    def getQuaternionFromAxisAngle(self, axis, angle):
        n = math.sqrt(sum(float(v) ** 2 for v in axis)); s = math.sin(angle / 2) / n
        return (axis[0] * s, axis[1] * s, axis[2] * s, math.cos(angle / 2))

    def getContactPoints(self, bodyA=-1, bodyB=-1, linkIndexA=-2, linkIndexB=-2, **k):
        out = []
        for (a, la, b, lb) in self.contacts:
            if bodyA >= 0 and bodyB >= 0:
                if (a, b) == (bodyA, bodyB): pass
                elif (b, a) == (bodyA, bodyB): a, la, b, lb = b, lb, a, la  # reported with the requested A first
                else: continue
            elif bodyA >= 0:
                if a == bodyA: pass
                elif b == bodyA: a, la, b, lb = b, lb, a, la
                else: continue
            out.append((0, a, b, la, lb, (0.0,) * 3, (0.0,) * 3, (0.0, 0.0, 1.0), 0.0, 0.0, 0.0, (0.0,) * 3, 0.0, (0.0,) * 3))
        return tuple(out)


def robot_joint_table(urdf):
    """getJointInfo tuples in Bullet's depth-first numbering."""
    from compile_model import parse_urdf
    links, order = parse_urdf(urdf)
    tmap = {'revolute': 0, 'continuous': 0, 'prismatic': 1, 'fixed': 4}
    out = []
    for i, (child, j, parent) in enumerate(order):
        lim, dyn = j.find('limit'), j.find('dynamics')
        g = lambda e, k: float(e.get(k, 0.0)) if e is not None else 0.0
        axis = tuple(float(v) for v in j.find('axis').get('xyz').split()) if j.find('axis') is not None else (0.0, 0.0, 0.0)
        out.append((i, j.get('name').encode(), tmap[j.get('type')], -1, -1, 0, g(dyn, 'damping'), g(dyn, 'friction'),
                    g(lim, 'lower'), g(lim, 'upper'), g(lim, 'effort'), g(lim, 'velocity'), child.encode(), axis,
                    (0.0,) * 3, (0.0, 0.0, 0.0, 1.0), parent))
    return out


def install_stand_ins(bullet):
    class Box:
        def __init__(self, low, high, dtype=np.float32, **k):
            self.low, self.high, self.dtype = np.asarray(low), np.asarray(high), np.dtype(dtype)
            self.shape = self.low.shape

    class Env:
        def reset(self, *, seed=None, options=None):
            return None

    registry = {}
    gym = types.ModuleType('gymnasium'); gym.Env = Env
    spaces = types.ModuleType('gymnasium.spaces'); spaces.Box = Box
    envs = types.ModuleType('gymnasium.envs')
    reg = types.ModuleType('gymnasium.envs.registration')
    reg.register = lambda id, entry_point=None, max_episode_steps=None, kwargs=None, **k: registry.__setitem__(
        id, dict(entry_point=entry_point, max_episode_steps=max_episode_steps, kwargs=kwargs or {}))
    gym.spaces, gym.envs, envs.registration = spaces, envs, reg
    pdata = types.ModuleType('pybullet_data'); pdata.getDataPath = lambda: '/nonexistent'
    sys.modules.update({'pybullet': bullet, 'pybullet_data': pdata, 'gymnasium': gym, 'gymnasium.spaces': spaces,
                        'gymnasium.envs': envs, 'gymnasium.envs.registration': reg, 'cv2': types.ModuleType('cv2')})
    return registry


@contextlib.contextmanager
def quiet():
    with contextlib.redirect_stdout(io.StringIO()):
        yield


# ----------------------------------------------------------------------------------------------------- scenarios
ENV_CLASSES = {
    'BallBalance': ('urgym.envs.env_ball_balance_v0', 'BallBalance'),
    'TwoBallsBalance': ('urgym.envs.env_two_balls_balance_v0', 'TwoBallsBalance'),
    'CubesPush': ('urgym.envs.env_cubes_push_v0', 'CubesPush'),
    'CubesGrasp': ('urgym.envs.env_cubes_grasp_v0', 'CubesGrasp'),
    'Golf': ('urgym.envs.env_golf_v0', 'Golf'),
    'CubesPushDiff': ('urgym.envs.env_cubes_push_diff_v0', 'CubesPush'),
    'GolfJoints': ('urgym.envs.env_golf_joints_v0', 'Golf'),
}
SCENE_OF = {'CubesPushDiff': 'CubesPush', 'GolfJoints': 'Golf'}   # variants live in their parents' worlds
# the 12 revolute joints in the oracle's DoF order = ascending pybullet joint index
DOF_JOINTS = [1, 2, 3, 4, 5, 6, 9, 11, 13, 14, 16, 18]


def make_env(name, bullet, **kw):
    import importlib
    bullet.clear()
    mod = importlib.import_module(ENV_CLASSES[name][0])
    with quiet():
        if name == 'CubesGrasp':
            env = getattr(mod, ENV_CLASSES[name][1])(render_mode='train')
        else:
            env = getattr(mod, ENV_CLASSES[name][1])(render_mode='train', **kw)
    return env


def body_names(name, env):
    """pybullet body id -> role name used in the fixture."""
    name = SCENE_OF.get(name, name)
    m = {env.robot.id: 'robot'}
    if name in ('BallBalance', 'TwoBallsBalance', 'CubesPush', 'Golf'):
        m[env.table_id] = 'table'
    m[env.planeID] = 'plane'
    if name == 'BallBalance':
        m[env.paddle_id] = 'free0'; m[env.ball_id] = 'free1'
    elif name == 'TwoBallsBalance':
        m[env.paddle_id] = 'free0'; m[env.ball1_id] = 'free1'; m[env.ball2_id] = 'free2'
    elif name == 'CubesPush':
        m[env.cubes[0]] = 'free0'; m[env.cubes[1]] = 'free1'
    elif name == 'CubesGrasp':
        m[env.cubes[0]] = 'free0'
    elif name == 'Golf':
        m[env.pitch_id] = 'pitch'; m[env.stick_id] = 'free0'; m[env.ball_id] = 'free1'
    return m


def script_kinematics(bullet, env, kin, q):
    """Load the oracle's kinematic answers for joint vector q into the scripted world."""
    rid = env.robot.id
    for d, j in enumerate(DOF_JOINTS):
        bullet.joint_pos[(rid, j)] = float(q[d])
    ee = tuple(kin['ee'][:3]), tuple(kin['ee'][3:])
    bullet.link_state[(rid, 7)] = (ee[0], ee[1], (0.0,) * 3, (0.0, 0.0, 0.0, 1.0), ee[0], ee[1])
    for link, key in ((11, 'link11'), (16, 'link16')):
        pos = tuple(kin[key])
        bullet.link_state[(rid, link)] = (pos, (0.0, 0.0, 0.0, 1.0), (0.0,) * 3, (0.0, 0.0, 0.0, 1.0), pos, (0.0, 0.0, 0.0, 1.0))
    # two different boxes whose union is centred on the oracle's gripper centre
    c = np.asarray(kin['gripper_center'])
    bullet.aabb[(rid, 11)] = (tuple(c - [0.05, 0.01, 0.02]), tuple(c + [0.01, 0.01, 0.01]))
    bullet.aabb[(rid, 16)] = (tuple(c - [0.01, 0.02, 0.01]), tuple(c + [0.05, 0.02, 0.02]))


def script_free_bodies(bullet, env, name, state, kin):
    ids = {v: k for k, v in body_names(name, env).items()}
    for b in range(3):
        key = 'free%d' % b
        if key in ids:
            p = state[38 + 13 * b: 38 + 13 * b + 13]
            bullet.bodies[ids[key]]['pos'] = tuple(float(v) for v in p[0:3])
            bullet.bodies[ids[key]]['orn'] = tuple(float(v) for v in p[3:7])
    if SCENE_OF.get(name, name) == 'Golf':
        foot = tuple(kin['club_foot'])
        bullet.link_state[(ids['free0'], 0)] = (foot, (0.0, 0.0, 0.0, 1.0), (0.0,) * 3, (0.0, 0.0, 0.0, 1.0), foot, (0.0, 0.0, 0.0, 1.0))
    return ids


def commands_since(bullet, mark):
    """Normalised command events issued since log index `mark`."""
    ev, log, i = [], bullet.log[mark:], 0
    while i < len(log):
        n, a, k = log[i]
        if n == 'resetJointState':
            vals = []
            while i < len(log) and log[i][0] == 'resetJointState':
                vals.append(log[i][1][1]); i += 1
            ev.append(dict(op='reset_arm', q=vals)); continue
        if n == 'stepSimulation':
            c = 0
            while i < len(log) and log[i][0] == 'stepSimulation':
                c += 1; i += 1
            ev.append(dict(op='ticks', n=c)); continue
        if n == 'calculateInverseKinematics':
            ev.append(dict(op='ik', link=a[0], pos=a[1], orn=a[2], nullspace=k.get('nullspace'), maxNumIterations=k.get('maxNumIterations')))
            arm = []
            j = i + 1
            while j < len(log) and log[j][0] == 'setJointMotorControl2' and log[j][1][0] in (1, 2, 3, 4, 5, 6):
                arm.append(log[j]); j += 1
            ev[-1]['arm_motor'] = [dict(joint=m[1][0], target=m[2].get('targetPosition'), force=m[2].get('force'),
                                        maxVelocity=m[2].get('maxVelocity')) for m in arm]
            i = j; continue
        if n == 'setJointMotorControl2':
            if a[0] == 9:
                ev.append(dict(op='gripper', target=k.get('targetPosition'), force=k.get('force'), maxVelocity=k.get('maxVelocity')))
                i += 1; continue
            arm = []
            while i < len(log) and log[i][0] == 'setJointMotorControl2' and log[i][1][0] in (1, 2, 3, 4, 5, 6):
                arm.append(log[i]); i += 1
            if arm:
                ev.append(dict(op='joints', targets=[m[2].get('targetPosition') for m in arm], force=[m[2].get('force') for m in arm],
                               maxVelocity=[m[2].get('maxVelocity') for m in arm]))
                continue
        if n == 'createMultiBody':
            ev.append(dict(op='spawn', body=k['body'], pos=k['basePosition'], orn=k['baseOrientation'], mass=k['baseMass'], shape=k['shape']))
        elif n == 'loadURDF':
            ev.append(dict(op='spawn', body=k['body'], pos=a[1], orn=a[2], urdf=a[0], globalScaling=k.get('globalScaling')))
        elif n == 'settle':
            ev.append(dict(op='settle'))
        elif n in ('removeBody', 'changeDynamics'):
            ev.append(dict(op=n, args=a, kw=k))
        i += 1
    return ev


def main():
    import oracle_lib as O
    O.build()
    urdf = os.path.join(REF, 'urgym', 'urdf', 'ur5_robotiq_85.urdf')
    bullet = ScriptedBullet(robot_joint_table(urdf))
    registry = install_stand_ins(bullet)
    sys.path.insert(0, REF)
    with quiet():
        import urgym  # noqa: F401  (runs urgym/envs/__init__.py -> register())
        from urgym.base import utilities as U
    G = {'_about': 'Generated by tests/golden/make_golden.py from the reference\'s own Python run against a scripted '
                   'pybullet stand-in. Physics (Bullet C++) is NOT pinned by this file.',
         'registry': registry}

    # ---- helper math (utilities.py)
    rng = np.random.default_rng(20240611)
    gdr = []
    for ts, tm in ((0.1, 0.5), (0.5, 2.0)):
        for v in list(np.linspace(0.0, 1.2 * tm, 25)) + [1e-12, ts, tm, tm + 1e-9, 3 * tm]:
            gdr.append([float(v), ts, tm, float(U.geometric_distance_reward(float(v), ts, tm))])
    qm, qn, qr = [], [], []
    for _ in range(12):
        a, b = rng.normal(size=4), rng.normal(size=4)
        a /= np.linalg.norm(a); b /= np.linalg.norm(b)
        qm.append([a.tolist(), b.tolist(), _jsonable(U.quaternion_multiply(a.tolist(), b.tolist()))])
        c = rng.normal(size=4) * 3
        qn.append([c.tolist(), _jsonable(U.normalize_quaternion(*c.tolist()))])
        ang, ax = float(rng.uniform(-3, 3)), rng.normal(size=3)
        ax /= np.linalg.norm(ax)
        qr.append([a.tolist(), ang, ax.tolist(), _jsonable(U.rotate_quaternion(a.tolist(), ang, ax.tolist()))])
    qr.append([[0.0, 0.0, 0.0, 1.0], math.pi / 2, [0, 1, 0], _jsonable(U.rotate_quaternion([0.0, 0.0, 0.0, 1.0], math.pi / 2, [0, 1, 0]))])
    G['utilities'] = dict(geometric_distance_reward=gdr, quaternion_multiply=qm, normalize_quaternion=qn, rotate_quaternion=qr)

    G['envs'] = {}
    for name in ENV_CLASSES:
        t = O.ENV_TYPES[name]
        E = {}
        env = make_env(name, bullet)
        E['spaces'] = dict(obs_low=_jsonable(env.observation_space.low), obs_high=_jsonable(env.observation_space.high),
                           obs_dtype=str(env.observation_space.dtype), act_low=_jsonable(env.action_space.low),
                           act_high=_jsonable(env.action_space.high), act_dtype=str(env.action_space.dtype))
        r = env.robot
        E['robot'] = dict(eef_id=r.eef_id, arm_num_dofs=r.arm_num_dofs, rest_pose=_jsonable(r.arm_rest_poses),
                          controllable_joints=_jsonable(r.controllable_joints), arm_joints=_jsonable(r.arm_controllable_joints),
                          gripper_range=_jsonable(r.gripper_range), mimic_parent=r.mimic_parent_id,
                          mimic_children=_jsonable(r.mimic_child_multiplier),
                          arm_lower=_jsonable(r.arm_lower_limits), arm_upper=_jsonable(r.arm_upper_limits),
                          max_force=[r.joints[j].maxForce for j in r.controllable_joints],
                          max_velocity=[r.joints[j].maxVelocity for j in r.controllable_joints],
                          control_method=env.control_method)
        E['construction'] = [e for e in bullet.log if e[0] in ('setGravity', 'loadURDF', 'createCollisionShape', 'createMultiBody',
                                                               'changeDynamics', 'createConstraint', 'changeConstraint')]
        # gripper opening length -> commanded finger_joint angle
        ga = []
        for pct in (0, 20, 50, 100):
            m = len(bullet.log); r.open_gripper(pct)
            ga.append(dict(call='open_gripper', percentage=pct, target=commands_since(bullet, m)[0]['target']))
            m = len(bullet.log); r.close_gripper(pct)
            ga.append(dict(call='close_gripper', percentage=pct, target=commands_since(bullet, m)[0]['target']))
        E['gripper'] = ga

        # ---- wait_until_stable(): scripted joint motion, count the stepSimulation calls it makes
        ws = []
        for label, deltas in (('still', [0.0]), ('decay', [0.1, 0.05, 0.006, 0.0051, 0.005, 0.00499, 0.0]),
                              ('spread', [0.012, 0.0061, 0.0059, 0.0]), ('never', [0.01] * 600)):
            rid = r.id
            for j in DOF_JOINTS:
                bullet.joint_pos[(rid, j)] = 0.0
            seq = list(deltas)

            def on_step(b, seq=seq, rid=rid):
                d = seq[min(b.nsteps - 1, len(seq) - 1)]
                for j in DOF_JOINTS:  # spread the L1 motion over all 12 controllable joints
                    b.joint_pos[(rid, j)] += d / 12 * (1 if b.nsteps % 2 else -1)
            bullet.nsteps, bullet.on_step = 0, on_step
            with quiet():
                ok = env.wait_until_stable()
            ws.append(dict(label=label, l1_motion_per_tick=deltas, ticks=bullet.nsteps, returned=bool(ok)))
            bullet.on_step = None
        E['wait_until_stable'] = ws

        # ---- reset(): command sequence for seeded draws, under two different scripted kinematics
        ob = O.OracleBatch(name, 1, seed=3)
        ob.reset()
        resets = []
        for k, (seed, dq) in enumerate(((11, 0.0), (12, 0.15), (13, -0.1))):
            st = ob.get_state()[0]
            st[0:6] += dq
            ob.set_state(st)
            kin = ob.kinematics()
            script_kinematics(bullet, env, kin, st[0:12])
            draws = []
            real_uniform = random.uniform

            def rec_uniform(a, b, draws=draws):
                v = real_uniform(a, b); draws.append([a, b, v, (v - a) / (b - a)]); return v
            random.seed(seed); random.uniform = rec_uniform
            env.wait_until_stable = lambda *a, **kk: bullet._rec('settle')
            mark = len(bullet.log)
            try:
                with quiet():
                    env.reset()
            finally:
                random.uniform = real_uniform
                del env.wait_until_stable
            ev2 = commands_since(bullet, mark)
            names = body_names(name, env)
            for e in ev2:
                if e['op'] == 'spawn':
                    e['role'] = names.get(e['body'], 'marker')
            resets.append(dict(seed=seed, draws=draws, kin={kk: _jsonable(v) for kk, v in kin.items()}, events=ev2,
                               vertical_quaternion=_jsonable(getattr(env, 'vertical_quaternion', None))))
        E['reset'] = resets

        # ---- step(): scripted states -> reference outputs
        E['step'] = step_cases(name, env, bullet, O)
        if name in SCENE_OF:   # the variants share their parents' predicates: a third of the cases covers what differs
            E['step'] = E['step'][::3]
        G['envs'][name] = E

    out = os.path.join(HERE, 'reference_env_logic.json')
    with open(out, 'w') as f:
        json.dump(G, f, indent=0, separators=(',', ':'))
    print('wrote', out, os.path.getsize(out), 'bytes;', {n: len(G['envs'][n]['step']) for n in G['envs']}, 'step cases')


def contact_menu(name):
    """Contact sets (role, link, role, link) worth exercising for each env's predicates."""
    name = SCENE_OF.get(name, name)
    R, none = 'robot', []
    if name in ('BallBalance', 'TwoBallsBalance'):
        return [none, [(R, 12, 'free0', -1), (R, 17, 'free0', -1), ('free1', -1, 'free0', -1)]]
    if name == 'CubesPush':
        return [none,
                [(R, 0, 'table', -1)],                                   # base link only: not a collision
                [(R, 3, 'table', -1)], [(R, 12, 'table', -1), (R, 12, 'free1', -1)],
                [(R, 12, 'free1', -1)], [('free1', -1, R, 9)], [(R, 18, 'free1', -1)], [(R, 8, 'free1', -1)], [(R, 6, 'free1', -1)],
                [(R, 12, 'free0', -1)], [('free0', -1, 'free1', -1), (R, 17, 'free1', -1)],
                [('free0', -1, 'table', -1), ('free1', -1, 'table', -1)]]
    if name == 'CubesGrasp':
        return [none, [(R, 12, 'free0', -1)], [(R, 17, 'free0', -1)], [(R, 12, 'free0', -1), (R, 17, 'free0', -1)],
                [(R, 11, 'free0', -1), (R, 16, 'free0', -1)], [(R, 13, 'free0', -1), (R, 18, 'free0', -1)],
                [(R, 10, 'free0', -1), (R, 15, 'free0', -1)], [(R, 12, 'free0', -1), (R, 13, 'free0', -1)],
                [(R, 12, 'plane', -1), (R, 17, 'plane', -1)], [('free0', -1, 'plane', -1), (R, 11, 'free0', -1), (R, 18, 'free0', -1)]]
    if name == 'Golf':
        hold = [(R, 12, 'free0', -1), (R, 17, 'free0', -1)]
        return [none, hold, [('free0', -1, R, 9)], [(R, 8, 'free0', -1)], [(R, 12, 'free0', 1)],
                hold + [('free0', 0, 'pitch', -1)], hold + [(R, 12, 'pitch', -1)], hold + [('free1', -1, 'free0', 0)],
                hold + [('free1', -1, 'free0', 0), ('free0', 0, 'pitch', -1)], hold + [('free1', -1, 'pitch', -1)],
                hold + [('free0', 0, 'table', -1)], [('free1', -1, 'free0', -1)]]
    raise KeyError(name)


def step_cases(full_name, env, bullet, O):
    name = SCENE_OF.get(full_name, full_name)
    rng = np.random.default_rng(sum(map(ord, full_name)))
    ob = O.OracleBatch(full_name, 1, seed=21)
    ob.reset()
    base = ob.get_state()[0].copy()
    ad = ob.ad
    lo, hi = np.asarray(env.action_space.low, float), np.asarray(env.action_space.high, float)
    cases = []
    # make sure the env object has its bodies (run one scripted reset)
    kin = ob.kinematics()
    script_kinematics(bullet, env, kin, base[0:12])
    with quiet():
        env.wait_until_stable = lambda *a, **k: True
        env.reset()
        del env.wait_until_stable

    def variants():
        """Yield (label, state) — realistic states from the oracle plus edits that reach every branch."""
        yield 'after_reset', base.copy()
        s = base.copy()
        for k in range(3):  # a few simulated steps for realistic joint / body configurations
            a = rng.uniform(lo, hi)
            ob.step(a, u=np.array([0.3]))
            s = ob.get_state()[0].copy()
            yield 'sim_step_%d' % k, s.copy()
        fb = lambda b: slice(38 + 13 * b, 41 + 13 * b)
        if name in ('BallBalance', 'TwoBallsBalance'):
            e = s.copy(); e[fb(1)] = e[fb(0)] + [0.01, -0.02, -0.03]; yield 'ball1_below_paddle', e
            e = s.copy(); e[fb(1)] = e[fb(0)] + [0.02, 0.03, 0.02]; yield 'ball1_above_paddle', e
            e = s.copy(); e[fb(1)] = e[fb(0)] + [0.0, 0.0, 0.0]; yield 'ball1_level_with_paddle', e
            e = s.copy(); e[41:45] = quat_from_euler([0.3, -0.2, 1.1]); yield 'paddle_tilted', e
            e = s.copy(); e[41:45] = quat_from_euler([0.1, 1.5707, 0.4]); yield 'paddle_near_gimbal', e
            if name == 'TwoBallsBalance':
                e = s.copy(); e[fb(2)] = e[fb(0)] + [0.01, 0.02, -0.01]; yield 'ball2_below_paddle', e
                e = s.copy(); e[fb(1)] = [1.2, -0.4, 1.0]; e[fb(2)] = [-0.3, -1.5, 1.0]; yield 'balls_far', e
                e = s.copy(); e[fb(1)] = [1.0, -1.0, 1.0]; yield 'ball_at_unit_distance', e
        elif name == 'CubesPush':
            e = s.copy(); e[fb(0)] = [0.0, -0.5, 0.02]; e[fb(1)] = [0.03, -0.53, 0.02]; yield 'cubes_closer_than_5cm', e
            e = s.copy(); e[fb(0)] = [0.0, -0.5, 0.02]; e[fb(1)] = [0.05, -0.5, 0.02]; yield 'cubes_exactly_5cm', e
            e = s.copy(); e[fb(0)] = [0.0, -0.5, 0.02]; e[fb(1)] = [0.0, -0.58, 0.02]; yield 'cubes_8cm', e
            e = s.copy(); e[fb(0)] = [0.0, -0.5, 0.02]; e[fb(1)] = [0.3, -0.9, 0.02]; yield 'cubes_50cm', e
            e = s.copy(); e[fb(0)] = [0.0, -0.5, 0.02]; e[fb(1)] = [0.6, -0.9, 0.02]; yield 'cubes_beyond_max', e
        elif name == 'CubesGrasp':
            kin = None
            for flags in (0, 4, 12):
                e = s.copy(); e[31] = (int(e[31]) & ~12) | flags; yield 'flags_%d' % flags, e
                ob.set_state(e); gc = ob.kinematics()['gripper_center']
                f = e.copy(); f[fb(0)] = gc + [0.01, 0.0, -0.02]; yield 'flags_%d_cube_in_gripper' % flags, f
                f = e.copy(); f[fb(0)] = gc + [0.0, 0.0, 0.05]; yield 'flags_%d_cube_at_5cm' % flags, f
                f = e.copy(); f[fb(0)] = [gc[0], gc[1], 0.25]; yield 'flags_%d_cube_raised' % flags, f
                f = e.copy(); f[fb(0)] = [gc[0], gc[1], 0.2]; yield 'flags_%d_cube_at_raise_height' % flags, f
                g = e.copy(); g[6:12] = [0.78, -0.78, 0.78, 0.78, -0.78, 0.78]  # gripper closed: opening < 0.5
                ob.set_state(g); gc2 = ob.kinematics()['gripper_center']
                g[fb(0)] = gc2 + [0.0, 0.01, 0.0]; yield 'flags_%d_closed_gripper_on_cube' % flags, g
        elif name == 'Golf':
            e = s.copy(); e[fb(1)] = [0.01, -0.62, 0.03]; yield 'ball_in_hole', e
            e = s.copy(); e[fb(1)] = [0.0, -0.65, 0.02]; yield 'ball_exactly_5cm_from_hole', e
            e = s.copy(); e[fb(1)] = [0.3, -0.9, -0.05]; yield 'ball_below_zero', e
            e = s.copy(); e[fb(1)] = [0.2, -0.8, 0.06]; yield 'ball_on_pitch', e
            ob.set_state(s); foot = ob.kinematics()['club_foot']
            e = s.copy(); e[fb(1)] = foot + [0.03, 0.02, 0.0]; yield 'ball_next_to_club', e

    menu = contact_menu(name)
    ids = None
    for label, st in variants():
        for ci, contacts in enumerate(menu):
            for reward_type in (('dense', 'sparse') if name in ('CubesPush', 'Golf') else ('dense',)):
                if reward_type == 'sparse' and ci % 3:
                    continue
                action = rng.uniform(lo, hi)
                if name == 'CubesGrasp':
                    mode = len(cases) % 4
                    if mode == 0: action[0:2] = [1.0, 0.0]
                    elif mode == 1: action[0:2] = [0.0, 1.0]
                    elif mode == 2: action[0:2] = [0.0, 0.0]
                    action[8] = [0.1, 0.9, 0.5, 0.49][len(cases) % 4]
                action = action.astype(np.float32).astype(np.float64)  # what a float32 action space delivers
                ob.set_state(st)
                kin = ob.kinematics()
                script_kinematics(bullet, env, kin, st[0:12])
                ids = script_free_bodies(bullet, env, name, st, kin)
                bullet.contacts = [(ids[a], la, ids[b], lb) for (a, la, b, lb) in contacts]
                if hasattr(env, 'reward_type'):
                    env.reward_type = reward_type
                if name == 'CubesGrasp':
                    fl = int(st[31])
                    env.subgoals_achieved = {'approached': bool(fl & 4), 'grasped': bool(fl & 8)}
                if name in ('CubesPush', 'Golf'):
                    env.vertical_quaternion = [float(v) for v in st[34:38]]
                seed = 1000 + len(cases)
                random.seed(seed); u = random.random(); random.seed(seed)
                mark = len(bullet.log)
                try:
                    with quiet():
                        obs, reward, terminated, truncated, info = env.step(action.copy())
                except UnboundLocalError as ex:  # Golf, reward_type='sparse', no floor hit and no touch: `success` is never
                    cases.append(dict(label=label, contacts=[list(c) for c in contacts], reward_type=reward_type,  # assigned
                                      state=_jsonable(st[:77]), action=_jsonable(action), u=u, raises=type(ex).__name__))
                    continue
                ev = [{k: v for k, v in e.items() if k not in ('nullspace', 'arm_motor')}
                      for e in commands_since(bullet, mark) if e['op'] in ('ik', 'joints', 'gripper')]
                case = dict(label=label, contacts=[list(c) for c in contacts], reward_type=reward_type, state=_jsonable(st[:77]),
                            action=_jsonable(action), u=u, commands=ev, obs=_jsonable(np.asarray(obs, float)), reward=float(reward),
                            terminated=bool(terminated), truncated=bool(truncated), info=_jsonable(info),
                            settle_ticks_in_scripted_world=sum(1 for e in bullet.log[mark:] if e[0] == 'stepSimulation'))
                if name == 'CubesGrasp':
                    case['subgoals_after'] = _jsonable(env.subgoals_achieved)
                cases.append(case)
    return cases


if __name__ == '__main__':
    main()
