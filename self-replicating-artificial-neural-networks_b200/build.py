"""Builds libserann_b200.so (CUDA kernels + C ABI) in-tree for sm_100a with nvcc.

    python -m serann_b200.build        # or: from serann_b200.build import build; build()

The library is rebuilt when any source under csrc/ or include/ is newer than it.
"""
from __future__ import annotations

import os
import pathlib
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = pathlib.Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / 'csrc'
LIB = PKG / 'libserann_b200.so'
SOURCES = ['srn_engine.cu', 'srn_ops.cu', 'srn_decode.cu']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC', '-Xcompiler', '-fvisibility=default', '--expt-relaxed-constexpr']


def _nvcc() -> str:
    nvcc = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(nvcc):
        raise RuntimeError('nvcc not found: libserann_b200.so cannot be built')
    return nvcc


def _stale() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = list(CSRC.glob('*')) + list((ROOT / 'include').glob('*.h'))
    return any(p.stat().st_mtime > t for p in deps if p.suffix in ('.cu', '.cuh', '.h'))


def build(force: bool = False, verbose: bool = False) -> pathlib.Path:
    if not force and not _stale():
        return LIB
    nvcc = _nvcc()
    objdir = PKG / 'build'
    objdir.mkdir(exist_ok=True)

    def compile_one(src: str) -> pathlib.Path:
        obj = objdir / (src + '.o')
        cmd = [nvcc, *NVCC_FLAGS, '-c', str(CSRC / src), '-o', str(obj)]
        if verbose:
            cmd.insert(1, '-Xptxas=-v')
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f'nvcc failed on {src}:\n{r.stdout}\n{r.stderr}')
        if verbose:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    tmp = LIB.with_suffix('.so.tmp')
    r = subprocess.run([nvcc, '-shared', '-o', str(tmp), *map(str, objs), '-lcudart'], capture_output=True, text=True)
    if r.returncode:
        raise RuntimeError(f'link failed:\n{r.stdout}\n{r.stderr}')
    os.replace(tmp, LIB)
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
