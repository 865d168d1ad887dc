"""A scripted CubesGrasp episode that goes through the physics: open -> above the cube -> down around it -> close -> raise.

Exercises the branches a random policy never reaches (env_cubes_grasp_v0.py:150-168, 206-233): 'approached' (gripper AABB
centre within 0.05 of the cube with the gripper open, +5), 'grasped' (contact on a link of {11,12,13} and of {16,17,18}, +5),
raised (grasped and cube z > 0.2, +10, terminated, is_success). Joint targets come from the oracle's IK iterated to
convergence on the oracle's own state (test infrastructure); the resulting ACTIONS are what a backend under test is fed.
"""
import numpy as np

import oracle_lib as O

VERTICAL = np.array([0.0, np.sin(np.pi / 4), 0.0, np.cos(np.pi / 4)])   # rotate_quaternion(identity, pi/2, y): gripper pointing down
RAISES = (0.06, 0.12, 0.18, 0.24, 0.30)


def _ik(q12, pos, iters=12):
    q = np.array(q12, float)
    for _ in range(iters):
        q[:6] = O.ik(q, np.concatenate([pos, VERTICAL]))
    return q[:6]


def _action(n, targets=None, grip=None):
    a = np.zeros((n, 9))
    if targets is not None:
        a[:, 0] = 1.0; a[:, 2:8] = targets          # selector weights (1, 0): always "move" (env_cubes_grasp_v0.py:127-130)
    else:
        a[:, 1] = 1.0; a[:, 8] = grip               # (0, 1): always "gripper"; bit < 0.5 opens, else closes
    return a.astype(np.float32).astype(np.float64)


def run(ob, n, on_step=None):
    """Drives the oracle batch `ob` (freshly reset, n envs) through the script. on_step(k, action, oracle_result) is called after
    every gym step. Returns the list of (action, oracle result)."""
    log = []

    def step(a):
        r = ob.step(a, u=np.full(n, 0.5))
        log.append((a, r))
        if on_step:
            on_step(len(log) - 1, a, r)
        return r

    r = step(_action(n, grip=0.0))
    for zoff in (0.30, 0.175):
        st = ob.get_state()
        cube = r['obs'][:, 10:13].copy()            # the cube is dropped from z = 0.1 at reset: it has landed after the first move
        tg = np.stack([_ik(st[i, :12], cube[i] + np.array([0, 0, zoff])) for i in range(n)])
        r = step(_action(n, targets=tg))
    r = step(_action(n, grip=1.0))
    for dz in RAISES:
        st = ob.get_state()
        tg = np.stack([_ik(st[i, :12], cube[i] + np.array([0, 0, 0.175 + dz])) for i in range(n)])
        r = step(_action(n, targets=tg))
        if r['terminated'].all():
            break
    return log
