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
LIB_CHECKS = PKG / 'libodk_checks.so'  # same sources with -DODK_CHECKS: device-side bounds assertions (csrc/odk_common.cuh)

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


def build(force: bool = False, verbose: bool = False, checks: bool = False) -> pathlib.Path:
    """One nvcc -c per source file, in parallel (objects under csrc/.obj/, git-ignored), then one link."""
    lib = LIB_CHECKS if checks else LIB
    if not checks and not force and not needs_build():
        return LIB
    import concurrent.futures as cf
    nvcc = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    obj_dir = PKG / 'csrc' / ('.obj_checks' if checks else '.obj')
    obj_dir.mkdir(exist_ok=True)
    hdr_time = max(p.stat().st_mtime for p in HDR + [pathlib.Path(__file__)])
    compile_flags = [f for f in NVCC_FLAGS if f != '-shared'] + (['-DODK_CHECKS=1'] if checks else [])

    def compile_one(src):
        obj = obj_dir / (src.stem + '.o')
        if not force and obj.exists() and obj.stat().st_mtime > max(src.stat().st_mtime, hdr_time):
            return obj, ''
        cmd = [nvcc] + compile_flags + (['-Xptxas', '-v'] if verbose else []) + ['-c', '-o', str(obj), str(src)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f'nvcc failed on {src.name}:\n' + res.stdout + res.stderr)
        return obj, res.stderr

    with cf.ThreadPoolExecutor(max_workers=len(SRC)) as pool:
        results = list(pool.map(compile_one, SRC))
    if verbose:
        for _, log in results:
            print(log)
    res = subprocess.run([nvcc, '-gencode', 'arch=compute_100a,code=sm_100a', '-shared', '-Xcompiler', '-fPIC', '-o', str(lib)]
                         + [str(o) for o, _ in results], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError('nvcc link failed:\n' + res.stdout + res.stderr)
    return lib


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv, checks='--checks' in sys.argv))
