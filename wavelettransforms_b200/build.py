"""Builds wavelettransforms_b200/lib/libwtprune.so in-tree with nvcc for sm_100a.

    python -m wavelettransforms_b200.build [--force]

-fmad=false: the DWT arithmetic must not be contracted into FMAs; the kernels also
use explicit __fmul_rn/__fadd_rn, the flag is the belt to those braces.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIBDIR = os.path.join(HERE, 'lib')
OBJDIR = os.path.join(HERE, 'build')
LIB = os.path.join(LIBDIR, 'libwtprune.so')

SOURCES = ['wt_fused_s7.cu', 'wt_fused_s5.cu', 'wt_fused_s3.cu', 'wt_fused_s1.cu', 'wt_small.cu', 'wt_small_fma.cu', 'wt_fused_flat.cu', 'wt_api.cu',
           'wt_select.cu', 'wt_generic.cu', 'wt_generic_fma.cu', 'wt_maskers.cu', 'wt_blocks.cu', 'wt_noalloc.cu']  # heaviest first
# the fma_taporder instantiations of the fused kernels (wt_fused_s<N>f.cu) double the build time: WT_BUILD_FMA=1 adds them
if os.environ.get('WT_BUILD_FMA'):
    SOURCES = ['wt_fused_s7f.cu', 'wt_fused_s5f.cu', 'wt_fused_s3f.cu', 'wt_fused_s1f.cu'] + SOURCES
    NVCC_EXTRA = ['-DWT_WITH_FMA_KERNELS']
else:
    NVCC_EXTRA = []
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17', '-fmad=false',
              '-diag-suppress', '177', '-Xcompiler', '-fPIC']


def _nvcc():
    for cand in (os.environ.get('NVCC'), '/usr/local/cuda/bin/nvcc', 'nvcc'):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return 'nvcc'


def _deps_mtime():
    m = 0.0
    for d in (CSRC, os.path.join(os.path.dirname(HERE), 'include')):
        for f in os.listdir(d):
            if f.endswith(('.cu', '.cuh', '.h', '.inc')):
                m = max(m, os.path.getmtime(os.path.join(d, f)))
    return m


def build_library(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= _deps_mtime():
        return LIB
    nvcc = _nvcc()
    hdr_m = max(os.path.getmtime(os.path.join(d, f))
                for d in (CSRC, os.path.join(os.path.dirname(HERE), 'include'))
                for f in os.listdir(d) if f.endswith(('.cuh', '.h', '.inc')))

    def compile_one(src):
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJDIR, src.replace('.cu', '.o'))
        if not force and os.path.exists(o) and os.path.getmtime(o) >= max(os.path.getmtime(s), hdr_m):
            return o
        cmd = [nvcc] + NVCC_FLAGS + NVCC_EXTRA + ['-c', s, '-o', o]
        if verbose:
            print(' '.join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f'nvcc failed for {src}:\n{r.stdout}\n{r.stderr}')
        return o

    with ThreadPoolExecutor(max_workers=min(os.cpu_count() or 4, len(SOURCES))) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc, '-shared', '-o', LIB] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a', '-lcudart']
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f'link failed:\n{r.stdout}\n{r.stderr}')
    return LIB


if __name__ == '__main__':
    print(build_library(force='--force' in sys.argv, verbose=True))
