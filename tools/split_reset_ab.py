"""Auto-reset inside the step launch (URGYM_SPLIT_RESET=0) against a reset-only launch after it (=1), DESIGN.md §4.5.
python tools/split_reset_ab.py ball4096 [steps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from urgym_b200.batch import EnvBatch  # noqa: E402


def run(env, n, K, W, split):
    os.environ['URGYM_SPLIT_RESET'] = str(int(split))
    dev = torch.device('cuda', 0)
    b = EnvBatch(env, n, device=0, seed=20261104)
    gen = torch.Generator(device=dev); gen.manual_seed(1000)
    acts = [bench.action_sampler(env, n, gen, torch, dev) for _ in range(K + W)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    b.reset()
    ms = 0.0
    for i in range(K + W):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        b.step(acts[i], autoreset=True)
        e1.record()
        flush.fill_(i & 255)
        torch.cuda.synchronize()
        if i >= W:
            ms += e0.elapsed_time(e1)
    st = b.stats()
    b.close()
    return n * K / (ms * 1e-3), st['ticks']


if __name__ == '__main__':
    wl = sys.argv[1] if len(sys.argv) > 1 else 'ball4096'
    K = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    env, n, _ = bench.WORKLOADS[wl]
    for split in (0, 1, 0, 1):
        v, ticks = run(env, n, K, 4, split)
        print(wl, 'split_reset', split, 'env-steps/s %.0f' % v, 'ticks', ticks, flush=True)
