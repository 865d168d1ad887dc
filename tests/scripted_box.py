"""A scripted solution of Box-v0 (env_box_v0.py's task: open the box, press the button, close the box) that drives any backend
with the env's own actions: grasp the handle bar, lift the door along its arc (rotating the wrist with it) past 1.9 rad, let go,
press the button with the closed gripper, push the door back over the vertical so that it falls shut. Used by the tests to
exercise the articulated box, its limits / motor and the finger-on-bar / finger-on-button contacts through the physics."""
import numpy as np

HINGE = np.array([0.075, 0.0, 0.1])
BAR = np.array([-0.055, 0.0, 0.15])          # handle bar h3 with the door closed


def _qmul(a, b):
    x1, y1, z1, w1 = a; x2, y2, z2, w2 = b
    return np.array([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                     w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2])


def run(step, n, max_steps=200):
    """step(actions[n,8]) -> dict(obs[n,7], reward[n], terminated[n]); the script is open-loop in everything but the
    end-effector pose it reads from the observation. Returns the list of per-step results."""
    log = []

    def act(dpos, dq=None, grip=0.085):
        a = np.zeros((n, 8)); a[:, :3] = np.clip(dpos, -0.05, 0.05)
        if dq is not None:
            a[:, 3:7] = np.clip(dq, -0.314, 0.314)
        a[:, 7] = grip
        a = a.astype(np.float32).astype(np.float64)
        r = step(a); log.append(r)
        return r

    def goto(r, target, k, grip):
        for _ in range(k):
            r = act(target - r['obs'][:, :3], grip=grip)
        return r
    r = act(np.zeros((n, 3)))
    r = goto(r, BAR + [0, 0, 0.30], 12, 0.085)          # above the bar, gripper open
    r = goto(r, BAR + [0, 0, 0.125], 8, 0.085)          # pads around the bar
    r = goto(r, BAR + [0, 0, 0.125], 3, 0.0)            # close on it
    ee0, q0 = r['obs'][:, :3].copy(), r['obs'][:, 3:7].copy()
    th = 0.0
    for _ in range(36):                                 # lift along the arc, wrist rotating with the door
        th = min(th + 0.06, 2.0)
        c, s = np.cos(th), np.sin(th)
        Ry = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
        tgt = HINGE + (ee0 - HINGE) @ Ry.T
        qd = np.stack([_qmul(np.array([0, np.sin(th / 2), 0, np.cos(th / 2)]), q) for q in q0])
        q = r['obs'][:, 3:7]
        qd = np.where((np.sum(qd * q, 1) < 0)[:, None], -qd, qd)
        r = act(tgt - r['obs'][:, :3], qd - q, grip=0.0)
    r = goto(r, r['obs'][:, :3].copy(), 2, 0.085)       # let go: the door rests open on its upper limit
    away = r['obs'][:, :3].copy() + 0.10 * np.array([np.sin(th), 0.0, np.cos(th)])   # back off ALONG the (rotated) fingers: they point to
    r = goto(r, away, 4, 0.085)                         # (-sin th, 0, -cos th) now, so that neither finger drags the bar; then up and over
    for _ in range(16):                                 # wrist back to vertical on the way up
        q = r['obs'][:, 3:7]; qd = np.where((np.sum(q0 * q, 1) < 0)[:, None], -q0, q0)
        r = act(np.array([0.35, 0.0, 0.55]) - r['obs'][:, :3], qd - q, grip=0.0)
    r = goto(r, np.array([0.0, 0.0, 0.55]), 9, 0.0)
    r = goto(r, np.array([0.0, 0.0, 0.32]), 6, 0.0)     # above the button, gripper closed
    for _ in range(8):
        r = act(np.array([0.0, 0.0, 0.05]) - r['obs'][:, :3], grip=0.0)
        if (r['reward'] > 0.5).all():
            break
    r = goto(r, np.array([0.0, 0.0, 0.55]), 7, 0.0)     # up again
    r = goto(r, np.array([0.33, 0.0, 0.55]), 9, 0.0)    # over to the far side of the open door
    r = goto(r, np.array([0.33, 0.0, 0.28]), 7, 0.0)
    for _ in range(18):                                 # sweep towards -x: pushes the door over the vertical, it falls shut
        r = act(np.array([-0.14, 0.0, 0.28]) - r["obs"][:, :3], grip=0.0)
        if r['terminated'].all():
            break
    for _ in range(10):
        if r['terminated'].all():
            break
        r = act(np.array([-0.14, 0.0, 0.45]) - r["obs"][:, :3], grip=0.0)
    return log
