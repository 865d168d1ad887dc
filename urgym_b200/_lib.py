"""ctypes binding of liburgym_b200.so (the C ABI in include/urgym_b200.h).

There is no CPU fallback: importing this module without the built library, or creating a batch without a CUDA
device, raises. PyTorch is used only for device memory and streams.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get('URGYM_SO') or os.path.join(HERE, 'liburgym_b200.so')   # URGYM_SO: A/B builds of the same ABI

ENV_TYPES = {'BallBalance': 0, 'TwoBallsBalance': 1, 'CubesPush': 2, 'CubesGrasp': 3, 'Golf': 4, 'CubesPushDiff': 5, 'GolfJoints': 6}
STATE_FLOATS = 96
MAX_CONTACTS = 16
NUM_LINKS = 19

_lib = None


class UrgymError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise UrgymError('liburgym_b200.so is not built (run `python -m urgym_b200.build` or __graft_entry__.build()); '
                         'urgym_b200 has no CPU path')
    L = C.CDLL(SO_PATH)
    vp, i32, u64 = C.c_void_p, C.c_int, C.c_uint64
    L.urgym_create.argtypes = [i32, i32, i32, u64, i32, i32, C.POINTER(vp)]
    L.urgym_destroy.argtypes = [vp]
    L.urgym_reset.argtypes = [vp, vp, vp, vp, vp, vp]
    L.urgym_step.argtypes = [vp] + [vp] * 10 + [i32, vp]
    L.urgym_step_host.argtypes = [vp] + [vp] * 10 + [i32]
    L.urgym_reset_host.argtypes = [vp, vp, vp, vp, vp]
    L.urgym_get_state.argtypes = [vp, vp, vp]
    L.urgym_set_state.argtypes = [vp, vp, vp]
    L.urgym_contacts.argtypes = [vp, vp, vp, vp]
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
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise UrgymError('urgym_b200 error %d: %s' % (rc, lib().urgym_last_error().decode()))


def shape_table(env_type):
    L = lib()
    n = L.urgym_num_shapes(env_type)
    out = []
    for i in range(n):
        b, l = C.c_int(), C.c_int()
        check(L.urgym_shape_info(env_type, i, C.byref(b), C.byref(l)))
        out.append((b.value, l.value))
    return out
