"""The drop-in boundary without a GPU: the shared library loads, exports every symbol include/urgym_b200.h declares,
answers the host-only queries, and FAILS LOUDLY (no CPU fallback) when asked to compute without a CUDA device."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'urgym_b200.h')


@pytest.fixture(scope='module')
def L(built):
    from urgym_b200 import _lib
    return _lib.lib()


def declared_functions():
    src = re.sub(r'/\*.*?\*/', '', open(HEADER).read(), flags=re.S)
    return sorted(set(re.findall(r'\b(urgym_[a-z_0-9]+)\s*\(', src)))


def test_header_declares_the_expected_entry_points():
    names = declared_functions()
    for must in ('urgym_create', 'urgym_destroy', 'urgym_reset', 'urgym_step', 'urgym_step_host', 'urgym_reset_host', 'urgym_get_state',
                 'urgym_set_state', 'urgym_contacts', 'urgym_tick', 'urgym_step_logic', 'urgym_fk', 'urgym_ik', 'urgym_gripper_center',
                 'urgym_stats', 'urgym_launch_count', 'urgym_last_error', 'urgym_obs_dim', 'urgym_act_dim', 'urgym_max_episode_steps'):
        assert must in names


def test_every_declared_symbol_is_exported(L):
    for name in declared_functions():
        assert hasattr(L, name), name
    out = subprocess.check_output(['nm', '-D', '--defined-only', os.path.join(ROOT, 'urgym_b200', 'liburgym_b200.so')], text=True)
    exported = set(re.findall(r' T (urgym_\w+)', out))
    assert exported == set(declared_functions())       # and nothing undeclared leaks out under the prefix


def test_header_compiles_as_plain_c(tmp_path):
    src = tmp_path / 'use.c'
    src.write_text('#include "urgym_b200.h"\nint main(void) { urgym_handle h = 0; (void)h; return URGYM_OK; }\n')
    subprocess.check_call(['gcc', '-std=c99', '-Wall', '-Werror', '-I', os.path.join(ROOT, 'include'), '-c', str(src), '-o', str(tmp_path / 'use.o')])


def test_no_torch_or_cxx_types_in_the_signatures():
    txt = re.sub(r'/\*.*?\*/', '', open(HEADER).read(), flags=re.S)   # comments may mention who owns the buffers
    assert 'torch' not in txt.lower() and 'std::' not in txt and 'at::' not in txt and 'Tensor' not in txt


def test_static_queries_work_without_a_device(L):
    dims = {0: (6, 3, 500), 1: (9, 3, 500), 2: (9, 3, 50), 3: (17, 9, 20), 4: (9, 3, 100), 5: (9, 3, 50), 6: (15, 6, 100), 7: (7, 8, 200)}
    for t, (od, ad, ms) in dims.items():
        assert (L.urgym_obs_dim(t), L.urgym_act_dim(t), L.urgym_max_episode_steps(t)) == (od, ad, ms)
        g, sm = C.c_int(), C.c_int()
        assert L.urgym_kernel_info(t, C.byref(g), C.byref(sm)) == 0
        assert g.value == 32 and 0 < sm.value <= (15.2, 15.2, 19.4, 15.2, 36, 19.4, 36, 19.6)[t] * 1024   # 14 (11 CubesPush / Box, 6 Golf scenes) resident envs per SM
    assert L.urgym_obs_dim(8) < 0
    assert L.urgym_real_bytes() == 4 and [L.urgym_num_pairs(t) for t in range(8)] == [73, 94, 73, 35, 134, 73, 134, 158]


def test_shape_tables_agree_with_the_oracle(L):
    import oracle_lib as O
    for name, t in O.ENV_TYPES.items():
        want = O.OracleBatch(name, 1).shape_table()
        assert L.urgym_num_shapes(t) == len(want)
        for i, (b, l) in enumerate(want):
            bb, ll = C.c_int(), C.c_int()
            assert L.urgym_shape_info(t, i, C.byref(bb), C.byref(ll)) == 0 and (bb.value, ll.value) == (b, l)


def _has_cuda():
    import torch
    return torch.cuda.is_available()


@pytest.mark.skipif(_has_cuda(), reason='checks the no-device behaviour')
def test_create_fails_loudly_without_cuda(L):
    h = C.c_void_p()
    rc = L.urgym_create(0, 4, 0, 0, 0, 0, C.byref(h))
    assert rc != 0 and not h.value
    msg = L.urgym_last_error().decode()
    assert 'CUDA' in msg and 'no CPU path' in msg
    from urgym_b200 import _lib
    from urgym_b200.batch import EnvBatch
    with pytest.raises(_lib.UrgymError):
        EnvBatch('BallBalance', 4)


def test_bad_arguments_are_rejected(L):
    assert L.urgym_create(9, 4, 0, 0, 0, 0, C.byref(C.c_void_p())) != 0
    assert L.urgym_create(0, 0, 0, 0, 0, 0, C.byref(C.c_void_p())) != 0
    assert L.urgym_step(None, *([None] * 10), 0, None) != 0


def test_parity_build_exports_the_same_entry_points():
    """liburgym_b200_f64.so = the same sources with real = double: same symbols, 8-byte scalars."""
    from urgym_b200 import _lib
    L64 = _lib.lib('f64')
    for name in declared_functions():
        assert hasattr(L64, name), name
    assert L64.urgym_real_bytes() == 8 and L64.urgym_obs_dim(3) == 17


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under urgym_b200/ may import, link or name it."""
    pkg = os.path.join(ROOT, 'urgym_b200')
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith(('.py', '.cu', '.cuh', '.cpp', '.h')):
                txt = open(os.path.join(dp, f), errors='ignore').read()
                assert 'urgo_' not in txt and 'oracle_lib' not in txt and 'liburgo' not in txt, os.path.join(dp, f)
    out = subprocess.check_output(['ldd', os.path.join(pkg, 'liburgym_b200.so')], text=True)
    assert 'urgo' not in out
