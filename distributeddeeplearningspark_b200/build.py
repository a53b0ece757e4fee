"""In-tree build of the native library: every ``csrc/*.cu|*.cpp`` -> ``_lib/libddls_b200.so`` for sm_100a.

    python -m distributeddeeplearningspark_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU; the resulting ``.so`` is git-ignored but ships to the GPU box.
"""
from __future__ import annotations

import argparse
import glob
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(ROOT, 'csrc')
OUT_DIR = os.path.join(PKG, '_lib')
OUT = os.path.join(OUT_DIR, 'libddls_b200.so')
OBJ_DIR = os.path.join(ROOT, 'build', 'obj')
ARCH = ['-gencode', 'arch=compute_100a,code=sm_100a']
NVCC_FLAGS = ARCH + ['-lineinfo', '-O3', '-std=c++17', '--use_fast_math', '-Xcompiler', '-fPIC',
                     '-Xcompiler', '-Wall', '-diag-suppress', '177,550']


def _nvcc() -> str:
    for c in (os.environ.get('NVCC'), shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if c and os.path.exists(c):
            return c
    raise RuntimeError('nvcc not found')


def _sources():
    return sorted(glob.glob(os.path.join(CSRC, '*.cu')) + glob.glob(os.path.join(CSRC, '*.cpp')))


def _digest(path: str) -> str:
    h = hashlib.sha1()
    for dep in [path] + sorted(glob.glob(os.path.join(CSRC, '*.cuh')) + glob.glob(os.path.join(CSRC, '*.h')) +
                               glob.glob(os.path.join(ROOT, 'include', '*.h'))):
        with open(dep, 'rb') as f:
            h.update(f.read())
    h.update(' '.join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _compile_one(nvcc: str, src: str, verbose: bool) -> str:
    os.makedirs(OBJ_DIR, exist_ok=True)
    obj = os.path.join(OBJ_DIR, os.path.basename(src) + '.o')
    stamp = obj + '.sha1'
    dig = _digest(src)
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
        return obj
    cmd = [nvcc] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + ['-x', 'cu', '-c', src, '-o', obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(' '.join(cmd) + '\n' + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f'nvcc failed on {src}')
    with open(stamp, 'w') as f:
        f.write(dig)
    return obj


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = _nvcc()
    srcs = _sources()
    if not srcs:
        raise RuntimeError(f'no sources under {CSRC}')
    if force and os.path.isdir(OBJ_DIR):
        shutil.rmtree(OBJ_DIR)
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(nvcc, s, verbose), srcs))
    os.makedirs(OUT_DIR, exist_ok=True)
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(OUT) or os.path.getmtime(OUT) < newest:
        cmd = [nvcc] + ARCH + ['-shared', '-o', OUT] + objs + ['-lcudart_static', '-ldl', '-lpthread', '-lrt']
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(' '.join(cmd) + '\n' + r.stdout + r.stderr)
            raise RuntimeError('link failed')
    return OUT


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--force', action='store_true')
    ap.add_argument('--verbose', action='store_true')
    a = ap.parse_args()
    print(build(a.force, a.verbose))
