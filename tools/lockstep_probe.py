"""Raw-tick and gym-step throughput for W warps per CTA (URGYM_WARPS), W warps starting every tick together."""
import os, sys, time
sys.path.insert(0, '.')
import numpy as np, torch
from urgym_b200.batch import EnvBatch
name, n = sys.argv[1], int(sys.argv[2])
gb = EnvBatch(name, n, seed=1)
gb.reset(); torch.cuda.synchronize()
s0 = gb.get_state().clone()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
gb.tick(20); torch.cuda.synchronize()
gb.set_state(s0)
e0.record(); gb.tick(100); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
chk = float(gb.get_state().double().sum())
rng = np.random.default_rng(0)
acts = [torch.tensor(rng.uniform(-0.1, 0.1, (n, gb.act_dim)) if name != 'CubesGrasp' else np.concatenate([rng.uniform(0, 1, (n, 2)), rng.uniform(-1, 1, (n, 6)), rng.uniform(0, 1, (n, 1))], 1), dtype=torch.float32, device='cuda') for _ in range(12)]
gb.set_state(s0)
for a in acts[:4]: gb.step(a)
torch.cuda.synchronize(); t0 = gb.stats()['ticks']
e0.record()
for a in acts[4:]: gb.step(a)
e1.record(); torch.cuda.synchronize()
ms2 = e0.elapsed_time(e1); t1 = gb.stats()['ticks']
print('W', os.environ.get('URGYM_WARPS', '1'), name, n, 'raw ticks/s %.3g' % (100 * n / ms * 1e3), 'checksum %.6f' % chk, '| steps: env-steps/s %.0f ticks/s %.3g' % (8 * n / ms2 * 1e3, (t1 - t0) / ms2 * 1e3), flush=True)
