"""Builds object_detector_b200/libodk.so (sm_100a) with nvcc, in-tree.

``python -m object_detector_b200.build`` or ``__graft_entry__.build()``.  nvcc cross-compiles on a
box without a GPU.  The built .so is git-ignored but travels with gpurun snapshots.
"""
import pathlib
import shutil
import subprocess
import sys

PKG = pathlib.Path(__file__).resolve().parent
SRC = sorted((PKG / 'csrc').glob('*.cu'))
HDR = sorted((PKG / 'csrc').glob('*.cuh')) + sorted((PKG.parent / 'include').glob('*.h'))
LIB = PKG / 'libodk.so'

NVCC_FLAGS = [
    '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
    '-fmad=false',  # bit-exactness: no implicit FMA contraction anywhere (SURVEY A-3)
    '-Xcompiler', '-fPIC', '-shared',
]


def needs_build() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    return any(p.stat().st_mtime > t for p in SRC + HDR + [pathlib.Path(__file__)])


def build(force: bool = False, verbose: bool = False) -> pathlib.Path:
    if not force and not needs_build():
        return LIB
    nvcc = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    cmd = [nvcc] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + ['-o', str(LIB)] + [str(s) for s in SRC]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError('nvcc failed:\n' + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
