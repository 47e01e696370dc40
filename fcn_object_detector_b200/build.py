"""In-tree build of libfcnod_b200.so (nvcc, sm_100a only).

    python -m fcn_object_detector_b200.build [--force]

The shared library is the product's only compute path; nothing here falls back
to CPU code.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OUT = os.path.join(HERE, 'libfcnod_b200.so')
NVCC = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')

COMMON = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
          '-Xcompiler', '-fPIC', '-Xcompiler', '-Wall', '-Xptxas', '-v']
# parity-critical translation units: no FMA contraction
SOURCES = [('net.cu', []), ('postnet.cu', ['-fmad=false']), ('layer.cu', ['-fmad=false']),
           ('preproc.cu', ['-fmad=false']), ('train_head.cu', ['-fmad=false']), ('photo.cu', ['-fmad=false']), ('train_net.cu', []), ('netdef.cpp', []), ('workspace.cpp', [])]


def _deps():
    out = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    out.append(os.path.join(HERE, '..', 'include', 'fcnod_b200.h'))
    return out


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(p) > t for p in _deps())


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    objdir = os.path.join(HERE, 'build')
    os.makedirs(objdir, exist_ok=True)

    # headers rebuild everything; a changed .cu / .cpp rebuilds only its own object
    hdrs = [p for p in _deps() if p.endswith(('.h', '.cuh'))] + [os.path.abspath(__file__)]
    hdr_t = max(os.path.getmtime(p) for p in hdrs)

    def compile_one(item):
        src, extra = item
        obj = os.path.join(objdir, src + '.o')
        if not force and os.path.exists(obj) and \
                os.path.getmtime(obj) > max(hdr_t, os.path.getmtime(os.path.join(CSRC, src))):
            return obj
        cmd = [NVCC] + COMMON + extra + ['-c', os.path.join(CSRC, src), '-o', obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError('nvcc failed for %s:\n%s\n%s' % (src, r.stdout, r.stderr))
        with open(os.path.join(objdir, src + '.ptxas.log'), 'w') as f:
            f.write(r.stderr)
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [NVCC, '-shared', '-o', OUT] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a']
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError('link failed:\n%s\n%s' % (r.stdout, r.stderr))
    return OUT


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
