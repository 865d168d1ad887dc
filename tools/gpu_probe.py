"""Throughput probe used while developing: steps/s and ticks/step per workload (not the bench)."""
import sys, time
sys.path.insert(0, '.')
import torch
from bench import action_sampler, WORKLOADS
from urgym_b200.batch import EnvBatch
names = sys.argv[1:] or list(WORKLOADS)
for w in names:
    env, n, _ = WORKLOADS[w]
    b = EnvBatch(env, n, seed=1)
    gen = torch.Generator(device='cuda'); gen.manual_seed(0)
    t = time.time(); b.reset(); torch.cuda.synchronize(); tr = time.time() - t
    for i in range(5): b.step(action_sampler(env, n, gen, torch, 'cuda'))
    torch.cuda.synchronize(); s0 = b.stats(); t = time.time(); K = 20
    for i in range(K): b.step(action_sampler(env, n, gen, torch, 'cuda'))
    torch.cuda.synchronize(); dt = time.time() - t; s1 = b.stats()
    print('%s: reset %.3fs; %d steps %.3fs -> %.0f env-steps/s, %.1f ticks/step, %.3g ticks/s, episodes %d succ %d mean_len %.1f mean_ret %.2f' % (
        w, tr, K, dt, n * K / dt, (s1['ticks'] - s0['ticks']) / (s1['env_steps'] - s0['env_steps']), (s1['ticks'] - s0['ticks']) / dt,
        s1['episodes'], s1['successes'], s1['length_sum'] / max(s1['episodes'], 1), s1['return_sum'] / max(s1['episodes'], 1)), flush=True)
    b.close()
