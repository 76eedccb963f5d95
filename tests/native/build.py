"""Build the native test harnesses (tests/native/*.cu -> tests/native/_built/*.so) for sm_100a.
Called by __graft_entry__.build(); nvcc cross-compiles without a GPU."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, '_built')
HARNESS = os.path.join(OUT, 'libblock_reduce_harness.so')
FUSED_HARNESS = os.path.join(OUT, 'libfused_reduce_harness.so')
SHARED_HARNESS = os.path.join(OUT, 'libshared_accum_harness.so')
_HEADERS = [os.path.join(ROOT, 'include', h) for h in ('dace_b200_block_reduce.cuh', 'dace_b200_fused_reduce.cuh',
                                                            'dace_b200_shared_accum.cuh')]


def _build_one(src: str, lib: str, force: bool) -> str:
    newest = max(os.path.getmtime(f) for f in [src] + _HEADERS)
    if not force and os.path.exists(lib) and os.path.getmtime(lib) >= newest:
        return lib
    os.makedirs(OUT, exist_ok=True)
    subprocess.check_call(['nvcc', '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
                           '-shared', '-Xcompiler', '-fPIC', '-I', os.path.join(ROOT, 'include'), src, '-o', lib])
    return lib


def build(force: bool = False) -> str:
    _build_one(os.path.join(HERE, 'fused_reduce_harness.cu'), FUSED_HARNESS, force)
    _build_one(os.path.join(HERE, 'shared_accum_harness.cu'), SHARED_HARNESS, force)
    return _build_one(os.path.join(HERE, 'block_reduce_harness.cu'), HARNESS, force)


if __name__ == '__main__':
    print(build(force=True))
