#!/usr/bin/env python
"""Headline benchmark: batched URGym env-steps per second on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload ball4096|two16384|push16384|grasp16384|golf16384]
    python bench.py --impl reference ...     # the CPU arm: the oracle port on the host cores (pybullet is absent)

One "step" is one gym step of every env of the batch: one launch of the fused env kernel (action -> IK -> 1..480
physics ticks -> reward / termination / observation -> in-kernel auto-reset of finished envs). Actions are synthetic
random-policy draws from each env's action space. `value` is timed on the device with inputs resident in HBM; `e2e`
goes through the C ABI's host-buffer entry (urgym_step_host: H2D of the actions, the kernel, D2H of obs / reward /
flags) and is timed with the host clock around the synchronous calls, over the same K steps.

The headline (`value`, `e2e`, `roofline`) is BASELINE.json config 1 (BallBalance x 4096 on one GPU). The same line carries a
`workloads` block with BASELINE configs 2-4, Golf and Box at 16 384 envs per GPU (value, ms per step, ticks per env-step, e2e),
measured right after the headline with fewer steps (--side-steps), so that the driver's BENCH / SCALE records hold every
config at every GPU count; `--no-side` skips them.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {  # name -> (env, envs per GPU, BASELINE.json config index)
    'ball4096': ('BallBalance', 4096, 1),
    'ball16384': ('BallBalance', 16384, None),
    'two16384': ('TwoBallsBalance', 16384, 2),
    'push16384': ('CubesPush', 16384, 3),
    'grasp16384': ('CubesGrasp', 16384, 4),
    'golf16384': ('Golf', 16384, None),
    'pushdiff16384': ('CubesPushDiff', 16384, None),
    'golfjoints16384': ('GolfJoints', 16384, None),
    'box16384': ('Box', 16384, None),
}
# algorithmic HBM bytes per env-step (SURVEY.md §8d / DESIGN.md §5): state once in, once out, action in, obs/reward/flags out
BYTES_PER_ENV_STEP = {'BallBalance': 562, 'TwoBallsBalance': 678, 'CubesPush': 606, 'CubesGrasp': 530, 'Golf': 606,
                      'CubesPushDiff': 606, 'GolfJoints': 642, 'Box': 586}
METRIC = 'env-steps/sec (batched, device-timed)'


def action_sampler(env, n, gen, torch, device):
    """Random policy over each env's action space (reference spaces, e.g. env_ball_balance_v0.py:52-54)."""
    if env == 'CubesGrasp':  # env_cubes_grasp_v0.py:74: [0,1]^2 x [-pi,pi]^6 x [0,1]
        a = torch.rand(n, 9, generator=gen, device=device)
        a[:, 2:8] = (a[:, 2:8] * 2 - 1) * 3.14159265
        return a
    if env == 'Box':         # env_box_v0.py:44: [-0.05, 0.05]^3 x [-pi/10, pi/10]^4 x [0, 0.085]
        a = torch.rand(n, 8, generator=gen, device=device) * 2 - 1
        a[:, :3] *= 0.05; a[:, 3:7] *= 0.314159265; a[:, 7] = (a[:, 7] + 1) * 0.0425
        return a
    return (torch.rand(n, 6 if env == 'GolfJoints' else 3, generator=gen, device=device) * 2 - 1) * 0.1


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)['hbm_gbs']), 'measured'
    return 6650.0, 'fallback'


def ncu_traffic(workload):
    """The committed ncu capture of this workload's env_kernel launches (profiles/ncu_traffic.json), or {}: DRAM bytes per gym
    step and issue-slot utilisation on the ELAPSED basis (sm__inst_executed.avg.per_cycle_elapsed / 4, time-weighted over the
    launches of a step: idle SMs count as idle)."""
    p = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get(workload) or {}
    return {}


class ClockSampler:
    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits', '-lms', '200'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            f = [x.strip() for x in r.split(',')]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[2 + k].lower().startswith('active'):
                    reasons.add(nm)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons), 'samples': len(sm)}


def calibrate_envs_per_thread(env, steps, min_seconds=6.0, lo=4, hi=256):
    """Envs per host thread such that `steps` gym steps take at least `min_seconds`: one thread steps 4 envs twice and the cost
    per env-step is scaled (every thread gets the same count, so the wall clock of the arm is that of one thread)."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    import oracle_lib as O
    b = O.OracleBatch(env, 4, seed=99)
    b.reset()
    rng = np.random.default_rng(0)
    a = np.concatenate([rng.uniform(0, 1, (4, 2)), rng.uniform(-np.pi, np.pi, (4, 6)), rng.uniform(0, 1, (4, 1))], 1) if env == 'CubesGrasp' else rng.uniform(-0.1, 0.1, (4, b.ad))
    t0 = time.perf_counter()
    for _ in range(3):
        b.step(a, autoreset=True)
    per = (time.perf_counter() - t0) / 12
    return int(min(hi, max(lo, -(-min_seconds // (per * max(steps, 1))))))


def cpu_arm(env, seconds, steps=None, warmup=0, envs_per_thread=4, seed=1234):
    """The oracle port on the host cores: one thread per core, each stepping its own envs (ctypes releases the GIL).

    If `steps` is given, every thread runs exactly warmup + steps gym steps on its envs and the timed part is the
    `steps`; otherwise it runs for about `seconds`. Returns (env_steps_per_s, cores, sample description, ms per step).
    """
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    import oracle_lib as O
    O.lib()
    cores = os.cpu_count() or 1
    res = [None] * cores
    start_evt = threading.Event()

    def work(t):
        rng = np.random.default_rng(seed + t)
        b = O.OracleBatch(env, envs_per_thread, seed=seed, offset=t * envs_per_thread)
        b.reset()

        def act():
            if env == 'CubesGrasp':
                return np.concatenate([rng.uniform(0, 1, (envs_per_thread, 2)), rng.uniform(-np.pi, np.pi, (envs_per_thread, 6)), rng.uniform(0, 1, (envs_per_thread, 1))], 1)
            if env == 'Box':
                return np.concatenate([rng.uniform(-0.05, 0.05, (envs_per_thread, 3)), rng.uniform(-np.pi / 10, np.pi / 10, (envs_per_thread, 4)), rng.uniform(0, 0.085, (envs_per_thread, 1))], 1)
            return rng.uniform(-0.1, 0.1, (envs_per_thread, b.ad))
        for _ in range(warmup):
            b.step(act(), autoreset=True)
        start_evt.wait()
        t0 = time.perf_counter(); n = 0
        while True:
            b.step(act(), autoreset=True); n += 1
            if steps is not None:
                if n >= steps:
                    break
            elif time.perf_counter() - t0 > seconds:
                break
        res[t] = (n * envs_per_thread, time.perf_counter() - t0, b.total_ticks())

    th = [threading.Thread(target=work, args=(t,)) for t in range(cores)]
    for x in th:
        x.start()
    time.sleep(0.2 if steps is None else 0.0)
    start_evt.set()
    for x in th:
        x.join()
    total = sum(r[0] for r in res); wall = max(r[1] for r in res)
    nsteps = steps if steps is not None else None
    desc = '%s, %d envs in all (%d host threads x %d envs each), random actions, auto-reset, %s, %.1f s wall (fp64 C++ restatement; pybullet is not installed on this box)' % (
        env, cores * envs_per_thread, cores, envs_per_thread, ('%d gym steps per env' % steps) if steps is not None else ('%.0f s' % seconds), wall)
    return total / wall, cores, desc, (wall / nsteps * 1e3 if nsteps else None), total


SIDE_WORKLOADS = ['two16384', 'push16384', 'grasp16384', 'golf16384', 'box16384']   # BASELINE configs 2-4, Golf and Box, 16 384 envs per GPU


def measure(env, n_envs, K, W, rank, world, local, torch, dist, with_e2e=True, with_clocks=False, seed=20261104):
    """W warm-up + K timed gym steps of `n_envs` envs of `env` on this rank's GPU: device-timed per-step CUDA events on the launch
    stream (max over ranks of the summed time), then the same K steps once more through the host-buffer entry (e2e)."""
    import ctypes as C
    from urgym_b200 import _lib
    from urgym_b200.batch import EnvBatch
    dev = torch.device('cuda', local)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize(dev)

    b = EnvBatch(env, n_envs, device=local, seed=seed, env_index_offset=rank * n_envs)
    gen = torch.Generator(device=dev); gen.manual_seed(1000 + rank)
    actions = [action_sampler(env, n_envs, gen, torch, dev) for _ in range(K + W)]
    # L2 flush between timed steps by READING a buffer twice the size of L2: the lines it leaves behind are clean, so nothing
    # of it is written back inside the next kernel's window (round 1 flushed by writing; ncu then charged the write-back of
    # the flush buffer to env_kernel)
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
    sink = torch.zeros(1, dtype=torch.int64, device=dev)
    b.reset()
    for i in range(W):
        b.step(actions[i]); sink += flush.view(torch.int64).sum()
    barrier()
    st0 = b.stats(); l0 = _lib.lib().urgym_launch_count()
    clocks = ClockSampler(local) if (with_clocks and rank == 0) else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    for i in range(K):
        ev[i][0].record()
        b.step(actions[W + i])
        ev[i][1].record()
        sink += flush.view(torch.int64).sum()
    barrier()
    l1 = _lib.lib().urgym_launch_count(); st1 = b.stats()
    ms_total = sum(a.elapsed_time(c) for a, c in ev)
    clk = clocks.stop() if clocks else None
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    out = {'value': world * n_envs * K / (ms_total * 1e-3), 'ms_per_step': ms_total / K, 'gpu_launches': int(l1 - l0), 'clocks': clk}
    gstats = torch.tensor([st1['ticks'] - st0['ticks'], st1['env_steps'] - st0['env_steps'],
                           st1['capacity_overflow_ticks'] - st0['capacity_overflow_ticks']], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(gstats)
    out['ticks_per_env_step'] = float(gstats[0].item() / max(gstats[1].item(), 1))
    out['physics_ticks_per_s'] = float(gstats[0].item() / (ms_total * 1e-3))
    out['capacity_overflow_tick_fraction'] = float(gstats[2].item() / max(gstats[0].item(), 1))
    # ---- e2e: host buffers through urgym_step_host (pinned), H2D + kernel(s) + D2H inside the timed region, all K steps
    if with_e2e:
        L = _lib.lib(); od, ad = b.obs_dim, b.act_dim
        h_act = [a.cpu().pin_memory() for a in actions[W:W + min(K, 16)]]
        h_obs = torch.empty(n_envs, od).pin_memory(); h_rew = torch.empty(n_envs).pin_memory()
        h_term = torch.empty(n_envs, dtype=torch.uint8).pin_memory(); h_trunc = torch.empty(n_envs, dtype=torch.uint8).pin_memory(); h_succ = torch.empty(n_envs, dtype=torch.uint8).pin_memory()
        p = lambda x: C.c_void_p(x.data_ptr())

        def estep(i):
            _lib.check(L.urgym_step_host(b.h, p(h_act[i % len(h_act)]), None, None, p(h_obs), p(h_rew), p(h_term), p(h_trunc), p(h_succ), None, None, 1))
        for i in range(min(3, W)):
            estep(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(K):
            estep(i)
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out['e2e'] = {'value': world * n_envs * K / float(t.item()), 'unit': 'env-steps/s', 'h2d_bytes_per_step': n_envs * ad * 4,
                      'd2h_bytes_per_step': n_envs * (od * 4 + 4 + 3), 'steps': K, 'timer': 'host clock around synchronous urgym_step_host calls'}
    b.close()
    del flush
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=100)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default='ball4096', choices=sorted(WORKLOADS))
    ap.add_argument('--envs', type=int, default=0, help='override envs per GPU')
    ap.add_argument('--cpu-seconds', type=float, default=12.0)
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-side', action='store_true', help='skip the `workloads` block (BASELINE configs 2-4 and Golf)')
    ap.add_argument('--side-steps', type=int, default=0, help='timed steps per side workload (default: max(6, steps // 8))')
    args = ap.parse_args()
    env, n_envs, cfg_idx = WORKLOADS[args.workload]
    if args.envs:
        n_envs = args.envs
    rank = int(os.environ.get('RANK', '0')); world = int(os.environ.get('WORLD_SIZE', '1')); local = int(os.environ.get('LOCAL_RANK', '0'))
    config = {'workload': 'URGym/%s-v0 batched %d envs per GPU, random policy, in-kernel auto-reset' % (env, n_envs),
              'baseline_config_index': cfg_idx, 'envs_per_gpu': n_envs, 'l2': 'L2 flushed between timed steps by reading a 256 MiB buffer (2 x L2)'}

    if args.impl == 'reference':
        # The reference's own implementation of the path is pybullet, which cannot be installed here or on the GPU box; this arm
        # times the fp64 C++ restatement (oracle/) on ALL host threads of this one host, whatever --gpus says, on a bounded
        # sample of the workload: `sample_envs` envs (not the GPU arm's batch) for `steps` gym steps, sized to run >= 6 s.
        if rank != 0:
            return
        ept = calibrate_envs_per_thread(env, args.steps + args.warmup)
        v, cores, desc, ms, total = cpu_arm(env, None, steps=args.steps, warmup=args.warmup, envs_per_thread=ept)
        config = dict(config, sample_envs=cores * ept, note='CPU arm: %d envs on %d host threads of ONE host%s; the GPU arm steps %d envs per GPU' % (
            cores * ept, cores, (' (the GPU arm uses %d GPUs; this arm does not scale with --gpus)' % args.gpus) if args.gpus > 1 else '', n_envs))
        line = {'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': 'env-steps/s', 'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
                'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
                'config': config, 'cpu_baseline': {'value': v, 'unit': 'env-steps/s', 'cores': cores, 'kind': 'port', 'sample': desc},
                'e2e': {'value': v, 'unit': 'env-steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}, 'gpu_launches': 0}
        print(json.dumps(line)); return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; urgym_b200 has no CPU path (use --impl reference for the CPU arm)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        # stdout carries exactly one JSON line: whatever native libraries print there (NCCL's "NCCL version ..." banner)
        # goes to stderr; the line itself is written to the saved descriptor
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        sys.stdout = os.fdopen(saved, 'w')
        dist.init_process_group('nccl', device_id=dev)

    K, W = args.steps, args.warmup
    main_res = measure(env, n_envs, K, W, rank, world, local, torch, dist, with_e2e=not args.no_e2e, with_clocks=True)
    side = {}
    if not args.no_side and args.workload == 'ball4096' and not args.envs:
        Ks = args.side_steps or max(6, K // 8)
        for w in SIDE_WORKLOADS:
            e2, n2, c2 = WORKLOADS[w]
            r = measure(e2, n2, Ks, max(3, min(W, 4)), rank, world, local, torch, dist, with_e2e=not args.no_e2e)
            side[w] = {'workload': 'URGym/%s-v0 batched %d envs per GPU' % (e2, n2), 'baseline_config_index': c2, 'value': r['value'], 'unit': 'env-steps/s',
                       'ms_per_step': r['ms_per_step'], 'steps': Ks, 'ticks_per_env_step': r['ticks_per_env_step'], 'physics_ticks_per_s': r['physics_ticks_per_s'],
                       'capacity_overflow_tick_fraction': r['capacity_overflow_tick_fraction'], 'e2e': r.get('e2e', {}).get('value'),
                       'hbm_frac': BYTES_PER_ENV_STEP[e2] * n2 / (r['ms_per_step'] * 1e-3) / 1e9 / peaks()[0]}
    if rank == 0:
        peak, how = peaks()
        ms_step = main_res['ms_per_step']
        achieved = BYTES_PER_ENV_STEP[env] * n_envs / (ms_step * 1e-3) / 1e9
        cap = ncu_traffic(args.workload if not args.envs else '')
        traffic = (cap['dram_bytes_read'] + cap['dram_bytes_write']) if cap else None
        line = {'metric': METRIC, 'value': main_res['value'], 'unit': 'env-steps/s', 'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': ms_step,
                'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic', 'config': config,
                'ticks_per_env_step': main_res['ticks_per_env_step'], 'physics_ticks_per_s': main_res['physics_ticks_per_s'],
                'capacity_overflow_tick_fraction': main_res['capacity_overflow_tick_fraction'],
                'gpu_launches': main_res['gpu_launches'], 'clocks': main_res['clocks'],
                'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                             'traffic': traffic, 'traffic_unit': 'bytes per gym step, summed over its env_kernel launches (ncu dram read + write)',
                             'algorithmic_bytes_per_launch': BYTES_PER_ENV_STEP[env] * n_envs, 'launches_per_step': main_res['gpu_launches'] / K,
                             'launch_note': 'one gym step = the env_kernel step grid, plus (ball scenes) the same kernel once more as a reset-only grid over the finished envs; achieved = bytes of the whole step / device time of the whole step',
                             'peak_source': how, 'kernel': 'env_kernel<%s>' % env,
                             'sm_issue_util_elapsed_pct': cap.get('issue_util_elapsed_pct'), 'issue_util_basis': 'sm__inst_executed.avg.per_cycle_elapsed / 4, time-weighted over the launches of a gym step',
                             'ncu_capture': cap.get('source'),
                             'note': 'bound by dependent fp32 / shuffle chains of the PGS sweeps, not by HBM; see DESIGN.md section 6 and profiles/'}}
        if 'e2e' in main_res:
            line['e2e'] = main_res['e2e']
        if side:
            line['workloads'] = side
        if world == 1 and not args.no_cpu:
            v, cores, desc, _, _ = cpu_arm(env, args.cpu_seconds)
            line['cpu_baseline'] = {'value': v, 'unit': 'env-steps/s', 'cores': cores, 'kind': 'port', 'sample': desc}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
