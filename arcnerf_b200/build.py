"""Builds arcnerf_b200/libarcb200.so (the C-ABI library, include/arcb200.h) with nvcc for sm_100a.

The library is built in-tree so that it travels with the repository snapshot to the GPU box; it is
git-ignored.  `python -m arcnerf_b200.build` or `__graft_entry__.build()` runs this.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OBJ_DIR = os.path.join(HERE, 'csrc', '_obj')
LIB_PATH = os.path.join(HERE, 'libarcb200.so')

ARCH = ['-gencode', 'arch=compute_100a,code=sm_100a']
COMMON = ['-O3', '-std=c++17', '-lineinfo', '-Xcompiler', '-fPIC', '--expt-relaxed-constexpr'] + os.environ.get('ARCB200_NVCC_EXTRA', '').split()

# Per-file extra flags.  The ray/occupancy kernels must not contract multiplies and adds (bit-exact
# parity of sample positions with the oracle); the rest follow the default (fmad on).
SOURCES = {
    'abi.cu': [],
    'march.cu': ['-fmad=false'],
    'occupancy.cu': ['-fmad=false'],
    'hashgrid.cu': [],
    'sh.cu': [],
    'composite.cu': [],
    'mlp_tc.cu': [],
    'optim.cu': [],
    'glue.cu': [],
    'noalloc.cu': [],
    'peer.cu': [],
    'zvals.cu': ['-fmad=false'],
    'volume_refresh.cu': [],
    'freq.cu': ['-fmad=false'],
    'probe.cu': [],
    'dense_chain.cu': [],
    'raygen.cu': ['-fmad=false'],
    'small_gemm.cu': [],
    'mcubes.cu': ['-fmad=false'],
    'surface.cu': ['-fmad=false'],
    'raster.cu': ['-fmad=false'],
}


def _nvcc():
    nvcc = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(nvcc):
        raise RuntimeError('nvcc not found: cannot build libarcb200.so')
    return nvcc


def _newer(path, deps):
    if not os.path.exists(path):
        return True
    t = os.path.getmtime(path)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    nvcc = _nvcc()
    os.makedirs(OBJ_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.cuh', '.h'))]
    headers.append(os.path.abspath(__file__))
    objs = []
    procs = []
    for src, extra in SOURCES.items():
        path = os.path.join(CSRC, src)
        if not os.path.exists(path):
            continue
        obj = os.path.join(OBJ_DIR, src.replace('.cu', '.o'))
        objs.append(obj)
        if force or _newer(obj, [path] + headers):
            cmd = [nvcc] + ARCH + COMMON + extra + ['-I', os.path.join(HERE, '..', 'include'), '-c', path, '-o', obj]
            if verbose:
                print(' '.join(cmd))
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write('nvcc failed for %s:\n%s\n' % (src, out))
        elif verbose and out.strip():
            print(out)
    if failed:
        raise RuntimeError('libarcb200 build failed')
    if force or procs or _newer(LIB_PATH, objs):
        cmd = [nvcc] + ARCH + ['-shared', '-o', LIB_PATH] + objs
        if verbose:
            print(' '.join(cmd))
        subprocess.check_call(cmd)
    return LIB_PATH


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose=True))
