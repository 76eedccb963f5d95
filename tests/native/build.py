"""Build the native test harnesses (tests/native/*.cu -> tests/native/_built/*.so) for sm_100a.
Called by __graft_entry__.build(); nvcc cross-compiles without a GPU."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, '_built')
HARNESS = os.path.join(OUT, 'libblock_reduce_harness.so')


def build(force: bool = False) -> str:
    src = os.path.join(HERE, 'block_reduce_harness.cu')
    hdr = os.path.join(ROOT, 'include', 'dace_b200_block_reduce.cuh')
    if not force and os.path.exists(HARNESS) and os.path.getmtime(HARNESS) >= max(os.path.getmtime(src),
                                                                                  os.path.getmtime(hdr)):
        return HARNESS
    os.makedirs(OUT, exist_ok=True)
    subprocess.check_call(['nvcc', '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
                           '-shared', '-Xcompiler', '-fPIC', '-I', os.path.join(ROOT, 'include'), src, '-o', HARNESS])
    return HARNESS


if __name__ == '__main__':
    print(build(force=True))
