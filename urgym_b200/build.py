"""Builds liburgym_b200.so (sm_100a) in-tree with nvcc. Cross-compiles without a GPU."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.environ.get('URGYM_SO') or os.path.join(HERE, 'liburgym_b200.so')
SRC = [os.path.join(HERE, 'csrc', f) for f in ('capi.cu', 'model_build.cpp')]


def _deps():
    d = [os.path.join(HERE, 'csrc', f) for f in os.listdir(os.path.join(HERE, 'csrc'))]
    return d + [os.path.join(HERE, '..', 'include', 'urgym_b200.h')]


def needs_build():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    return any(os.path.getmtime(d) > t for d in _deps())


def build(force=False, verbose=False):
    if not force and not needs_build():
        return SO
    nvcc = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
    cmd = [nvcc, '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
           '-Xcompiler', '-fPIC', '-shared', '-o', SO] + SRC
    if verbose:
        cmd += ['-Xptxas', '-v']
    cmd += os.environ.get('URGYM_NVCC_FLAGS', '').split()
    subprocess.check_call(cmd)
    return SO


if __name__ == '__main__':
    build(force='--force' in sys.argv, verbose='-v' in sys.argv)
    print(SO)
