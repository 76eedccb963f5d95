"""Which kernels would a reduction launch?  No GPU needed: asks the planner of the C ABI for its trace
(`dace_b200_reduce_nd_plan`) and the scratch it wants (`dace_b200_reduce_nd_workspace_bytes`).

    python tools/plan_of.py 1024,500,500 --axes 0,2 --strides 262144,512,1 [--dtype float32]
    python tools/plan_of.py 16,60,60,4000 --strides 16777216,262144,4096,1 --all-masks
"""
import argparse, ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dace_b200 import _lib, api


def dense(shape):
    st = [1] * len(shape)
    for i in range(len(shape) - 2, -1, -1):
        st[i] = st[i + 1] * shape[i + 1]
    return st


def plan(lib, dt, shape, strides, mask):
    n = len(shape)
    buf = ctypes.create_string_buffer(256)
    need = lib.dace_b200_reduce_nd_plan(dt, dt, n, (ctypes.c_int64 * n)(*shape), (ctypes.c_int64 * n)(*strides), mask, buf, 256)
    return buf.value.decode(), need - lib.dace_b200_workspace_bytes()


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('shape')
    ap.add_argument('--strides', default=None, help='input strides in elements (default: C-contiguous)')
    ap.add_argument('--axes', default=None, help='reduced axes (default: all)')
    ap.add_argument('--all-masks', action='store_true')
    ap.add_argument('--dtype', default='float32')
    a = ap.parse_args()
    shape = [int(x) for x in a.shape.split(',')]
    strides = [int(x) for x in a.strides.split(',')] if a.strides else dense(shape)
    dt = api.dtype_code(getattr(torch, a.dtype))
    lib = _lib.load()
    if a.all_masks:
        masks = range(1, 1 << len(shape))
    else:
        axes = [int(x) for x in a.axes.split(',')] if a.axes else range(len(shape))
        masks = [sum(1 << x for x in axes)]
    for m in masks:
        p, extra = plan(lib, dt, shape, strides, m)
        axes = tuple(i for i in range(len(shape)) if m >> i & 1)
        print('%-14s %-40s scratch beyond the base: %d bytes' % (axes, p, extra))
