"""In-tree build of libstgcn_b200.so (nvcc, sm_100a only).

`python -m stgcn_b200._build` or `__graft_entry__.build()`.  The .so lands next
to this file so it travels with the repo snapshot to the GPU box; it is
git-ignored.  nvcc cross-compiles without a GPU.
"""
import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OBJ = os.path.join(CSRC, 'build')
LIB = os.path.join(HERE, 'libstgcn_b200.so')
SOURCES = ['block.cu', 'bn_kernels.cu', 'agg_kernels.cu', 'gemm_simt.cu', 'head_kernels.cu', 'umma_gemm.cu',
           'umma_probe.cu', 'timing.cu', 'umma_wgrad.cu', 'umma_agg.cu', 'umma_spatial.cu', 'train_kernels.cu', 'input_kernels.cu', 'train_prep.cu']
NVCC_FLAGS = ['-std=c++17', '-O3', '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo',
              '-Xcompiler', '-fPIC', '--expt-relaxed-constexpr']
# development builds (e.g. STGCN_NVCC_EXTRA=-DSTGCN_SP_TRACE, umma_spatial.cu): part of the freshness digest, so a library built
# with extra flags is stale for a process that does not set the same variable
NVCC_FLAGS += os.environ.get('STGCN_NVCC_EXTRA', '').split()


def _nvcc():
    for c in (os.environ.get('NVCC'), '/usr/local/cuda/bin/nvcc', 'nvcc'):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return 'nvcc'


def have_nvcc():
    import shutil
    c = _nvcc()
    return os.path.exists(c) if os.path.isabs(c) else shutil.which(c) is not None


def _digest(paths):
    # file NAMES and contents only: the repository sits under a different absolute path on the GPU box, and a digest over
    # absolute paths made every process there consider the shipped library stale and rebuild it (8 ranks at once)
    h = hashlib.sha256()
    for p in sorted(paths, key=os.path.basename):
        h.update(os.path.basename(p).encode())
        with open(p, 'rb') as f:
            h.update(f.read())
    h.update(' '.join(NVCC_FLAGS).encode())
    return h.hexdigest()


def sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.cuh', '.h'))]
    hs.append(os.path.join(os.path.dirname(HERE), 'include', 'stgcn_b200.h'))
    return hs


def is_fresh():
    stamp = LIB + '.stamp'
    if not (os.path.exists(LIB) and os.path.exists(stamp)):
        return False
    with open(stamp) as f:
        return f.read().strip() == _digest(sources() + headers())


def build(force=False, verbose=False):
    """Compile every .cu for sm_100a and link the shared library. Returns its path.  Safe against concurrent callers (the
    ranks of one torchrun job): one process builds under a file lock, the others wait and find the result fresh; the library is
    linked to a temporary name and moved into place atomically."""
    if not force and is_fresh():
        return LIB
    import fcntl
    os.makedirs(OBJ, exist_ok=True)
    with open(os.path.join(OBJ, '.build.lock'), 'w') as lockf:
        fcntl.flock(lockf, fcntl.LOCK_EX)
        try:
            if not force and is_fresh():          # somebody else built it while we waited
                return LIB
            return _build_locked(verbose)
        finally:
            fcntl.flock(lockf, fcntl.LOCK_UN)


def _build_locked(verbose):
    nvcc = _nvcc()

    def compile_one(src):
        obj = os.path.join(OBJ, os.path.basename(src)[:-3] + '.o')
        cmd = [nvcc] + NVCC_FLAGS + ['-c', src, '-o', obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError('nvcc failed for %s:\n%s\n%s' % (src, r.stdout, r.stderr))
        if verbose and (r.stdout or r.stderr):
            print(r.stdout, r.stderr)
        return obj

    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, sources()))
    tmp = LIB + '.tmp.%d' % os.getpid()
    cmd = [nvcc, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', tmp] + objs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError('link failed:\n%s\n%s' % (r.stdout, r.stderr))
    os.replace(tmp, LIB)                        # atomic: a process loading the library never sees a half-written file
    with open(LIB + '.stamp', 'w') as f:
        f.write(_digest(sources() + headers()))
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose=True))
