"""Per-phase cycle breakdown of the tick (needs a build with URGYM_NVCC_FLAGS=-DUG_PHASE_TIMING)."""
import ctypes as C, sys
sys.path.insert(0, '.')
import torch
from bench import action_sampler, WORKLOADS
from urgym_b200 import _lib
from urgym_b200.batch import EnvBatch
L = _lib.lib()
names = ['fk', 'collide', 'dynamics', 'rows+pass1', 'T build', 'sweeps', 'integrate', '-']
for w in sys.argv[1:] or ['ball4096']:
    env, n, _ = WORKLOADS[w]
    b = EnvBatch(env, n, seed=1); gen = torch.Generator(device='cuda'); gen.manual_seed(0)
    b.reset()
    for i in range(5): b.step(action_sampler(env, n, gen, torch, 'cuda'))
    out = (C.c_ulonglong * 8)(); L.urgym_phase_cycles(out); s0 = b.stats()
    for i in range(10): b.step(action_sampler(env, n, gen, torch, 'cuda'))
    rc = L.urgym_phase_cycles(out); s1 = b.stats()
    ticks = s1['ticks'] - s0['ticks']; tot = sum(out)
    print(w, 'rc', rc, 'ticks', ticks, 'cycles/tick %.0f' % (tot / max(ticks, 1)))
    for k in range(7): print('  %-12s %8.0f cycles/tick  %5.1f%%' % (names[k], out[k] / max(ticks, 1), 100.0 * out[k] / max(tot, 1)))
    b.close()
