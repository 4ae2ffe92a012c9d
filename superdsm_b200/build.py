"""In-tree build of the CUDA library: ``python -m superdsm_b200.build`` (nvcc, sm_100a only)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, 'libsuperdsm_b200.so')
SOURCES = [os.path.join(HERE, 'csrc', name) for name in ('sdsm.cu', 'common.cuh', 'pack.cuh', 'grid.cuh', 'eval.cuh', 'ctrl.cuh', 'sched.cuh', 'image.cuh', 'logistic.cuh', 'preprocess.cuh')]
HEADER = os.path.join(os.path.dirname(HERE), 'include', 'superdsm_b200.h')


def needs_build():
    if not os.path.isfile(LIB_PATH): return True
    built = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(src) > built for src in SOURCES + [HEADER])


# -fmad=false: floating-point contraction is never left to the compiler (multiply-adds that should fuse are written as
# fma()): two evaluations of the energy at the same point are then bit-identical whatever code surrounds them, which the
# line searches of the solver rely on (csrc/eval.cuh).
def build(force=False, verbose=False):
    if not force and not needs_build(): return LIB_PATH
    nvcc = os.environ.get('NVCC', 'nvcc')
    cmd = [nvcc, '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17', '-fmad=false',
           '-Xcompiler', '-fPIC', '-shared', '-o', LIB_PATH, SOURCES[0]]
    cmd[1:1] = os.environ.get('SDSM_NVCC_FLAGS', '').split()      # e.g. -DSDSM_GRID_TRACE (phase timing of k_grid)
    if verbose: cmd[1:1] = ['-Xptxas', '-v']
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError('nvcc failed:\n' + res.stdout + res.stderr)
    if verbose: print(res.stderr)
    return LIB_PATH


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
