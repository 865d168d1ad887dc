"""Builds the C-ABI libraries (sm_100a) in-tree with nvcc. Cross-compiles without a GPU.

  liburgym_b200.so      the product: fp32 kernels
  liburgym_b200_f64.so  the parity build: the same sources with `real` = double (-DUG_F64), same entry points with
                        urgym_real = double (include/urgym_b200.h); used by the tests that compare integer outputs with
                        the fp64 oracle
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.environ.get('URGYM_SO') or os.path.join(HERE, 'liburgym_b200.so')
SO_F64 = os.environ.get('URGYM_SO_F64') or os.path.join(HERE, 'liburgym_b200_f64.so')
SRC = [os.path.join(HERE, 'csrc', f) for f in ('capi.cu', 'model_build.cpp')]


def _deps():
    d = [os.path.join(HERE, 'csrc', f) for f in os.listdir(os.path.join(HERE, 'csrc'))]
    return d + [os.path.join(HERE, '..', 'include', 'urgym_b200.h')]


def needs_build(so=None):
    so = so or SO
    if not os.path.exists(so):
        return True
    t = os.path.getmtime(so)
    return any(os.path.getmtime(d) > t for d in _deps())


def _compile(so, defines, verbose):
    nvcc = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
    cmd = [nvcc, '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
           '-Xcompiler', '-fPIC', '-shared', '-o', so] + defines + SRC
    if verbose:
        cmd += ['-Xptxas', '-v']
    cmd += os.environ.get('URGYM_NVCC_FLAGS', '').split()
    subprocess.check_call(cmd)


def build(force=False, verbose=False, f64=True):
    procs = []
    if force or needs_build(SO):
        _compile(SO, [], verbose)
    if f64 and (force or needs_build(SO_F64)):
        _compile(SO_F64, ['-DUG_F64'], verbose)
    return SO


if __name__ == '__main__':
    build(force='--force' in sys.argv, verbose='-v' in sys.argv, f64='--no-f64' not in sys.argv)
    print(SO)
