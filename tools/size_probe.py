"""Ad-hoc probe: device time vs input size for the full-sum and row-sum kernels (L2 flushed before each call)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
fb = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
add = 'lambda a, b: a + b'
def t(fn, k=20):
    for _ in range(3): fn()
    tot = 0.0
    for _ in range(k):
        fb.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / k * 1e3
for lg in (20, 22, 24, 25, 26, 27, 28):
    x = torch.rand(1 << lg, device=dev); out = torch.zeros(1, device=dev)
    us = t(lambda: db.reduce(add, x, out, axis=None, identity=0))
    print('full sum 2^%d f32 (%4d MiB): %8.2f us  %6.0f GB/s' % (lg, (4 << lg) >> 20, us, (4 << lg) / us / 1e3))
    del x
for rows, cols in ((4096, 4096), (16384, 1024), (1024, 16384), (65536, 256), (256, 65536), (8192, 8192)):
    a = torch.rand(rows, cols, device=dev); out = torch.zeros(rows, device=dev)
    us = t(lambda: db.reduce(add, a, out, axis=1, identity=0))
    print('row sum [%d, %d] f32 (%4d MiB): %8.2f us  %6.0f GB/s' % (rows, cols, rows * cols * 4 >> 20, us, rows * cols * 4 / us / 1e3))
    out2 = torch.zeros(cols, device=dev)
    us = t(lambda: db.reduce(add, a, out2, axis=0, identity=0))
    print('col sum [%d, %d] f32 (%4d MiB): %8.2f us  %6.0f GB/s' % (rows, cols, rows * cols * 4 >> 20, us, rows * cols * 4 / us / 1e3))
    del a
# empty-kernel floor
e = torch.zeros(1, device=dev)
print('torch tiny kernel floor: %.2f us' % t(lambda: e.add_(1)))
