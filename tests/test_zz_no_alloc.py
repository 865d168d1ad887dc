"""The allocation-free entry point (include/urgym_b200.h: urgym_step_static) against urgym_step. B200 only.

Collected last on purpose (the file name sorts after every other test file).
The entry has no counterpart in the reference; it was added for callers that cannot
call cudaMalloc. Bar: BIT-EXACT against the handle-based urgym_step from the same state records, the bar the suite uses for
two entries into the same kernel (test_host_entry_equals_device_entry), on two calls in a row (the second call runs on the
scratch contents the first one left behind).
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _buffers(n, od, dev, dt=torch.float32):
    return dict(state=torch.empty(n, 96, device=dev, dtype=dt), obs=torch.empty(n, od, device=dev, dtype=dt),
                reward=torch.empty(n, device=dev, dtype=dt), terminated=torch.empty(n, dtype=torch.uint8, device=dev),
                truncated=torch.empty(n, dtype=torch.uint8, device=dev), success=torch.empty(n, dtype=torch.uint8, device=dev),
                final_obs=torch.empty(n, od, device=dev, dtype=dt), nticks=torch.empty(n, 2, dtype=torch.int32, device=dev),
                stats=torch.empty(8, dtype=torch.float64, device=dev))


@pytest.mark.parametrize('name,n', [('BallBalance', 96), ('CubesGrasp', 64)])
def test_static_entry_equals_handle_entry_on_two_calls_in_a_row(built, name, n, monkeypatch):
    if not torch.cuda.is_available():
        pytest.fail('no CUDA device: the gpu tests must run on the B200 box')
    from urgym_b200.batch import EnvBatch, step_static
    if name == 'BallBalance':      # batches this small restart inside the step launch by default: ask for the two-launch form,
        monkeypatch.setenv('URGYM_SPLIT_RESET', '1')   # whose finished-env flags live in the static scratch
    seed, off = 11, 5
    gb = EnvBatch(name, n, seed=seed, env_index_offset=off)
    gb.reset()
    if name == 'BallBalance':      # make some episodes end inside the two steps so that the auto-reset (second launch) runs too
        s = gb.get_state(); s[::3, 32] = gb.max_episode_steps - 1; gb.set_state(s)
    dev = gb.device
    out = _buffers(n, gb.obs_dim, dev)
    rng = np.random.default_rng(4)
    state = gb.get_state().clone()
    for call in range(2):
        if name == 'CubesGrasp':
            a = np.concatenate([rng.uniform(0, 1, (n, 2)), rng.uniform(-np.pi, np.pi, (n, 6)), rng.uniform(0, 1, (n, 1))], 1)
        else:
            a = rng.uniform(-0.1, 0.1, (n, gb.act_dim))
        a = torch.tensor(a, dtype=torch.float32, device=dev)
        u = torch.tensor(rng.uniform(0, 1, n), dtype=torch.float32, device=dev)
        before = gb.stats()
        gb.step(a, u=u, autoreset=True)
        after = gb.stats()
        # poison the caller-owned outputs: everything compared below must have been written by this call
        for k, t in out.items():
            t.fill_(77)
        step_static(name, state, a, out, u=u, seed=seed, env_index_offset=off, autoreset=True)
        torch.cuda.synchronize()
        assert torch.equal(out['state'], gb.get_state()), (name, call)
        assert torch.equal(out['obs'], gb.obs) and torch.equal(out['reward'], gb.reward), (name, call)
        assert torch.equal(out['final_obs'], gb.final_obs) and torch.equal(out['nticks'], gb.nticks), (name, call)
        for k in ('terminated', 'truncated', 'success'):
            assert torch.equal(out[k], getattr(gb, k)), (name, call, k)
        # the call's own counters: episodes, env steps and physics ticks of this step only (cleared at the start of each call)
        st = out['stats'].cpu().numpy()
        assert st[4] == n and st[0] == after['episodes'] - before['episodes'] and st[5] == after['ticks'] - before['ticks'], (st, before, after)
        if name == 'BallBalance':
            assert st[0] > 0                      # resets did happen, through the scratch's finished-env flags
        state = out['state'].clone()
    gb.close()


def test_static_entry_rejects_batches_beyond_its_scratch(built):
    if not torch.cuda.is_available():
        pytest.fail('no CUDA device: the gpu tests must run on the B200 box')
    from urgym_b200 import _lib
    L = _lib.lib()
    n = _lib.STATIC_MAX_ENVS + 1
    one = torch.zeros(8, device='cuda')
    p = C.c_void_p(one.data_ptr())
    rc = L.urgym_step_static(0, 0, n, 0, 0, 0, p, p, None, None, p, p, p, p, p, p, None, None, None, 0, None)
    assert rc == -1                                # URGYM_E_ARG, before anything is launched
    msg = L.urgym_last_error().decode()
    assert 'URGYM_STATIC_MAX_ENVS' in msg and str(n) in msg
