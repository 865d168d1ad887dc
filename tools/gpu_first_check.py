import sys, os, time
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import numpy as np, torch
import oracle_lib as O
from urgym_b200.batch import EnvBatch, forward_kinematics, inverse_kinematics, gripper_center
rng = np.random.default_rng(0)
q = rng.uniform(-1.5, 1.5, (64, 12)); q[:, 6:] = rng.uniform(0, 0.8, (64, 1)) * np.array([1, -1, 1, 1, -1, 1])
fk_g = forward_kinematics(q).cpu().numpy().astype(np.float64)
fk_o = np.stack([O.fk(x) for x in q])
print('FK pos err', np.abs(fk_g[..., :3] - fk_o[..., :3]).max(), 'quat err', np.minimum(np.abs(fk_g[..., 3:] - fk_o[..., 3:]), np.abs(fk_g[..., 3:] + fk_o[..., 3:])).max(), 'signed', np.abs(fk_g[..., 3:] - fk_o[..., 3:]).max())
tg = fk_o[:, 7].copy(); tg[:, :3] += rng.uniform(-0.1, 0.1, (64, 3))
ik_g = inverse_kinematics(q, tg).cpu().numpy(); ik_o = np.stack([O.ik(q[i], tg[i]) for i in range(64)])
print('IK err', np.abs(ik_g - ik_o).max())
gc_g = gripper_center(q).cpu().numpy(); gc_o = np.stack([O.gripper_center(x) for x in q])
print('GC err', np.abs(gc_g - gc_o).max())
for name in ['BallBalance', 'TwoBallsBalance', 'CubesPush', 'CubesGrasp', 'Golf']:
    N = 8
    ob = O.OracleBatch(name, N, seed=3)
    gb = EnvBatch(name, N, seed=3)
    t = time.time(); oo, ont = ob.reset(); to = time.time() - t
    t = time.time(); go = gb.reset().cpu().numpy(); torch.cuda.synchronize(); tg_ = time.time() - t
    print(name, 'reset ticks oracle', ont, 'gpu', gb.reset_ticks.cpu().numpy(), 'obs err', np.abs(go - oo).max(), 'time o %.3f g %.3f' % (to, tg_))
    # identical initial state, single tick
    s0 = ob.get_state(); gb.set_state(torch.tensor(s0, dtype=torch.float32))
    s0f = gb.get_state().cpu().numpy().astype(np.float64); ob.set_state(s0f)
    ob.tick(1); gb.tick(1)
    s1o = ob.get_state(); s1g = gb.get_state().cpu().numpy()
    d = np.abs(s1g - s1o); print('  1-tick state err max', d.max(), 'argmax slot', np.unravel_index(d.argmax(), d.shape))
    ob.set_state(s0f); gb.set_state(torch.tensor(s0f, dtype=torch.float32))
    for k in range(5):
        a = rng.uniform(-0.1, 0.1, (N, ob.ad)) if name != 'CubesGrasp' else np.concatenate([rng.uniform(0, 1, (N, 2)), rng.uniform(-1, 1, (N, 6)), rng.uniform(0, 1, (N, 1))], 1)
        u = rng.uniform(0, 1, N)
        ro = ob.step(a, u=u); gb.step(a, u=u, autoreset=False)
        print('  step', k, 'nticks o', ro['nticks'][:, 0], 'g', gb.nticks[:, 0].cpu().numpy(), 'obs err %.2e' % np.abs(gb.obs.cpu().numpy() - ro['obs']).max(),
              'rew err %.2e' % np.abs(gb.reward.cpu().numpy() - ro['reward']).max(), 'term', ro['terminated'].astype(int), gb.terminated.cpu().numpy())
    print('  stats', gb.stats())
