"""Build libkgs.so in-tree with nvcc for sm_100a (no torch involved)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libkgs.so')
SOURCES = ['kgs.cu']
HEADERS = sorted(f for f in os.listdir(CSRC) if f.endswith(('.cuh', '.inc', '.h')))
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-shared', '-Xcompiler', '-fPIC']


def _stale():
    if not os.path.exists(LIB):
        return True
    built = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    deps.append(os.path.join(HERE, '..', 'include', 'kgs.h'))
    return any(os.path.exists(d) and os.path.getmtime(d) > built for d in deps)


def build_lib(force=False, verbose=False):
    """Compile csrc/kgs.cu -> kube_gpu_stats_b200/libkgs.so.  Returns the library path."""
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get('NVCC', 'nvcc')
    cmd = [nvcc] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + \
        ['-o', LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout + proc.stderr)
        raise RuntimeError('nvcc failed building libkgs.so')
    if verbose:
        sys.stderr.write(proc.stderr)
    return LIB


if __name__ == '__main__':
    print(build_lib(force='--force' in sys.argv, verbose='-v' in sys.argv))
