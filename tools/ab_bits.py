"""Bit-level A/B of prebuilt libraries with the same ABI: same seeds, same actions, digest of everything a rollout returns.

    python tools/ab_bits.py gpurun_ab/base.so urgym_b200/liburgym_b200.so:URGYM_RESET_GROUP=16 [workload ...]

Each variant (library, optionally with environment settings after a colon, comma-separated) runs in its own process;
a kernel change that claims "same operations, same order" must print identical digests.
"""
import hashlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child(workload, steps):
    sys.path.insert(0, ROOT)
    import torch
    from bench import WORKLOADS, action_sampler
    from urgym_b200.batch import EnvBatch
    env, n, _ = WORKLOADS[workload]
    n = min(n, 2048)
    b = EnvBatch(env, n, seed=11)
    gen = torch.Generator(device='cuda'); gen.manual_seed(5)
    h = hashlib.sha256()
    h.update(b.reset().cpu().numpy().tobytes()); h.update(b.reset_ticks.cpu().numpy().tobytes())
    for _ in range(steps):
        a = action_sampler(env, n, gen, torch, 'cuda')
        b.step(a)
        for t in (b.obs, b.reward, b.terminated, b.truncated, b.success, b.final_obs, b.nticks):
            h.update(t.cpu().numpy().tobytes())
    st = b.stats()
    print('%s ticks %d digest %s' % (workload, st['ticks'], h.hexdigest()[:24]))
    b.close()


if __name__ == '__main__':
    if sys.argv[1] == '--child':
        child(sys.argv[2], int(sys.argv[3]))
        sys.exit(0)
    libs = [a for a in sys.argv[1:] if '.so' in a]
    wls = [a for a in sys.argv[1:] if '.so' not in a] or ['ball4096', 'two16384', 'push16384', 'grasp16384', 'golf16384']
    bad = 0
    for w in wls:
        outs = []
        for so in libs:
            path, _, extra = so.partition(':')
            env = dict(os.environ, URGYM_SO=os.path.abspath(path))
            env.update(kv.split('=', 1) for kv in extra.split(',') if kv)
            r = subprocess.run([sys.executable, __file__, '--child', w, '12'], env=env, capture_output=True, text=True)
            outs.append(r.stdout.strip().splitlines()[-1] if r.returncode == 0 and r.stdout.strip() else 'FAILED ' + r.stderr[-300:])
        same = len(set(outs)) == 1
        bad += not same
        print(('SAME ' if same else 'DIFF ') + ' | '.join(outs))
    sys.exit(1 if bad else 0)
