"""Latency of the serial tick chains: device time of reset() (130-190 ticks per env, nothing to overlap with) and of one gym
step without auto-reset, for batches from one warp per SM up to the benchmark's size. With a -DUG_PHASE_TIMING library
(URGYM_SO=gpurun_ab/phase.so) also the per-phase warp cycles of those ticks.

    python tools/reset_latency.py [env ...]
"""
import ctypes as C
import sys
sys.path.insert(0, '.')
import torch
from bench import action_sampler
from urgym_b200 import _lib
from urgym_b200.batch import EnvBatch
L = _lib.lib()
names = ['fk', 'collide', 'dynamics', 'rows+pass1', 'T build', 'sweeps', 'integrate', '-']


def timed(fn):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record(); fn(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b)


def phases(label, ticks):
    out = (C.c_ulonglong * 8)()
    if L.urgym_phase_cycles(out) != 0:
        return
    tot = sum(out)
    print('    %s: %.0f cycles/tick:' % (label, tot / max(ticks, 1)), ', '.join('%s %.0f' % (names[k], out[k] / max(ticks, 1)) for k in range(7)))


for env in sys.argv[1:] or ['BallBalance', 'CubesGrasp']:
    for n in (148, 148 * 13, 4096, 16384):
        b = EnvBatch(env, n, seed=1); gen = torch.Generator(device='cuda'); gen.manual_seed(0)
        b.reset(); out = (C.c_ulonglong * 8)(); L.urgym_phase_cycles(out)
        s0 = b.stats(); ms = timed(b.reset); s1 = b.stats()
        tk = (s1['ticks'] - s0['ticks']) / n
        print('%s n=%d reset: %.2f ms, %.1f ticks per env -> %.1f us per tick of the chain' % (env, n, ms, tk, 1e3 * ms / tk))
        phases('reset', s1['ticks'] - s0['ticks'])
        for i in range(3):
            b.step(action_sampler(env, n, gen, torch, 'cuda'), autoreset=False)
        L.urgym_phase_cycles(out)
        a = action_sampler(env, n, gen, torch, 'cuda')
        s0 = b.stats(); ms = timed(lambda: b.step(a, autoreset=False)); s1 = b.stats()
        tk = (s1['ticks'] - s0['ticks']) / n
        mx = int(b.nticks[:, 0].max())
        print('%s n=%d step: %.2f ms, %.1f ticks per env (max %d) -> %.1f us per tick of the longest chain' % (env, n, ms, tk, mx, 1e3 * ms / mx))
        phases('step', s1['ticks'] - s0['ticks'])
        b.close()
