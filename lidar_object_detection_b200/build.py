"""Build liblidarfuse.so in-tree with nvcc for sm_100a (no JIT cache: the built
.so travels to the GPU box with the repository snapshot)."""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, '_lib', 'liblidarfuse.so')
SOURCES = ['ctx.cu', 'project.cu', 'erode.cu', 'fuse.cu', 'boxes2d.cu', 'resize.cu', 'colour.cu', 'csvfmt.cu', 'packhost.cu', 'points.cu', 'compact.cu', 'synth.cu', 'publish.cu']
HEADERS = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith('.cuh')) + \
    [os.path.join(ROOT, 'include', 'lidarfuse.h')]


def _nvcc():
    for cand in (os.environ.get('NVCC'), shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.isfile(cand):
            return cand
    raise RuntimeError('nvcc not found')


def needs_build():
    if not os.path.isfile(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + HEADERS
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo ... -> _lib/liblidarfuse.so"""
    if not force and not needs_build():
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    cmd = [_nvcc(), '-O3', '-std=c++17', '-lineinfo',
           '-gencode', 'arch=compute_100a,code=sm_100a',
           '-Xcompiler', '-fPIC,-O3', '-shared',
           '-I', os.path.join(ROOT, 'include'), '-I', CSRC,
           '-DLF_BUILDING=1']
    if verbose:
        cmd += ['-Xptxas', '-v']
    cmd += [os.path.join(CSRC, s) for s in SOURCES] + ['-o', LIB + '.tmp']
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError('nvcc failed:\n' + r.stdout + r.stderr)
    os.replace(LIB + '.tmp', LIB)
    if verbose:
        sys.stderr.write(r.stderr)
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
