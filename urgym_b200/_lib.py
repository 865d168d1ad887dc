"""ctypes binding of liburgym_b200.so (the C ABI in include/urgym_b200.h).

There is no CPU fallback: importing this module without the built library, or creating a batch without a CUDA
device, raises. PyTorch is used only for device memory and streams.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get('URGYM_SO') or os.path.join(HERE, 'liburgym_b200.so')   # URGYM_SO: A/B builds of the same ABI
# the parity build: same sources and entry points with urgym_real = double (include/urgym_b200.h, urgym_b200/build.py)
SO_PATH_F64 = os.environ.get('URGYM_SO_F64') or os.path.join(HERE, 'liburgym_b200_f64.so')

ENV_TYPES = {'BallBalance': 0, 'TwoBallsBalance': 1, 'CubesPush': 2, 'CubesGrasp': 3, 'Golf': 4, 'CubesPushDiff': 5, 'GolfJoints': 6, 'Box': 7}
STATE_FLOATS = 96
MAX_CONTACTS = 32      # width of the manifold-point export (include/urgym_b200.h: URGYM_MAX_CONTACTS)
PAIR_WORDS = 5         # uint32 words of the per-pair contact mask (URGYM_PAIR_WORDS)
NUM_LINKS = 19
STATIC_MAX_ENVS = 16384   # URGYM_STATIC_MAX_ENVS: largest batch of the allocation-free entry (urgym_step_static)

_libs = {}


class UrgymError(RuntimeError):
    pass


def lib(precision='f32'):
    """The product library (precision 'f32') or the fp64 parity build ('f64')."""
    if precision in _libs:
        return _libs[precision]
    path = {'f32': SO_PATH, 'f64': SO_PATH_F64}[precision]
    if not os.path.exists(path):
        raise UrgymError('%s is not built (run `python -m urgym_b200.build` or __graft_entry__.build()); '
                         'urgym_b200 has no CPU path' % os.path.basename(path))
    L = C.CDLL(path)
    vp, i32, u64 = C.c_void_p, C.c_int, C.c_uint64
    L.urgym_create.argtypes = [i32, i32, i32, u64, i32, i32, C.POINTER(vp)]
    L.urgym_destroy.argtypes = [vp]
    L.urgym_reset.argtypes = [vp, vp, vp, vp, vp, vp]
    L.urgym_step.argtypes = [vp] + [vp] * 10 + [i32, vp]
    L.urgym_step_static.argtypes = [i32, i32, i32, u64, i32, i32] + [vp] * 13 + [i32, vp]
    L.urgym_step_host.argtypes = [vp] + [vp] * 10 + [i32]
    L.urgym_reset_host.argtypes = [vp, vp, vp, vp, vp]
    L.urgym_get_state.argtypes = [vp, vp, vp]
    L.urgym_set_state.argtypes = [vp, vp, vp]
    L.urgym_contacts.argtypes = [vp, vp, vp, vp]
    L.urgym_contact_pairs.argtypes = [vp, vp, vp]
    L.urgym_pair_info.argtypes = [i32, i32, C.POINTER(i32), C.POINTER(i32)]
    L.urgym_tick.argtypes = [vp, i32, vp]
    L.urgym_step_logic.argtypes = [vp] + [vp] * 10
    L.urgym_fk.argtypes = [i32, i32, vp, vp, vp]
    L.urgym_ik.argtypes = [i32, i32, vp, vp, vp, vp]
    L.urgym_gripper_center.argtypes = [i32, i32, vp, vp, vp]
    L.urgym_stats.argtypes = [vp, C.POINTER(C.c_double)]
    L.urgym_shape_info.argtypes = [i32, i32, C.POINTER(i32), C.POINTER(i32)]
    L.urgym_kernel_info.argtypes = [i32, C.POINTER(i32), C.POINTER(i32)]
    L.urgym_launch_count.restype = C.c_longlong
    L.urgym_last_error.restype = C.c_char_p
    if L.urgym_real_bytes() != {'f32': 4, 'f64': 8}[precision]:
        raise UrgymError('%s was built for another scalar width than %s' % (path, precision))
    L.precision = precision
    _libs[precision] = L
    return L


def check(rc, L=None):
    if rc != 0:
        raise UrgymError('urgym_b200 error %d: %s' % (rc, (L or lib()).urgym_last_error().decode()))


def shape_table(env_type):
    L = lib()
    n = L.urgym_num_shapes(env_type)
    out = []
    for i in range(n):
        b, l = C.c_int(), C.c_int()
        check(L.urgym_shape_info(env_type, i, C.byref(b), C.byref(l)))
        out.append((b.value, l.value))
    return out


def pair_table(env_type, precision='f32'):
    """[(shape_a, shape_b)] of the scene's candidate pairs, in the bit order of urgym_contact_pairs."""
    L = lib(precision)
    out = []
    for i in range(L.urgym_num_pairs(env_type)):
        a, b = C.c_int(), C.c_int()
        check(L.urgym_pair_info(env_type, i, C.byref(a), C.byref(b)), L)
        out.append((a.value, b.value))
    return out
