"""One small call into every kernel family of the library, integer data, results compared bit for bit with torch.
Sizes are the smallest that still reach each kernel's dispatch threshold (many need >= 2^22 elements).  Written as
the driver for compute-sanitizer; that tool is closed on this GPU pool ("runs under it have left GPUs needing a
reset"), so it serves as a quick all-kernels check instead (~5 s): python tools/all_kernels_check.py [--big]"""
import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dace_b200 as db

dev = torch.device('cuda:0')
g = torch.Generator(device=dev); g.manual_seed(1)
ADD, MAX, MIN = 'lambda a, b: a + b', 'lambda a, b: max(a, b)', 'lambda a, b: min(a, b)'
done = []


def ints(shape, dt=torch.int32, lo=-50, hi=50):
    return torch.randint(lo, hi, shape, device=dev, dtype=dt, generator=g)


def red(name, x, axes, op=ADD, ident=0):
    out = db.reduce(op, x, axis=axes, identity=ident)
    if op == ADD:
        ref = x.sum(axes, dtype=x.dtype) if axes is not None else x.sum(dtype=x.dtype)
    else:
        ref = x.amax(axes) if axes is not None else x.max()
    assert torch.equal(out.reshape(ref.shape), ref), name
    done.append(name)


big = '--big' in sys.argv
# full reductions: split kernel (clusters + tickets), ragged length; the 1024-thread variant from 256 MiB
red('full 2^22+5', ints(((1 << 22) + 5, )), None)
if big:
    red('full 2^26 (big variant)', ints((1 << 26, )), None)
# rows
red('rows group [4096, 4096]', ints((4096, 4096)), (1, ))
red('rows group odd [5000, 1001]', ints((5000, 1001)), (1, ))
red('rows block [70, 40000]', ints((70, 40000)), (1, ))
red('rows split [3, 2^20]', ints((3, 1 << 20)), (1, ))
red('rows tiny [130000, 33]', ints((130000, 33)), (1, ))
red('rows tiny [45000, 100] max', ints((45000, 100)), (1, ), MAX, -1000)
red('rows tiny bf16-sized [140000, 33] i16', ints((140000, 33), torch.int16), (1, ))
buf = ints((70000 * 64 + 4, ))
red('rows tiny padded pitch [70000, 64] off by one', buf[1:1 + 70000 * 64].view(70000, 64), (1, ))
red('rows lanes-per-row [70000, 64]', buf[:70000 * 64].view(70000, 64), (1, ))
# columns
red('cols [65536, 256]', ints((65536, 256)), (0, ))
red('cols wide [4096, 8192]', ints((4096, 8192)), (0, ))
red('cols few rows [7, 600004]', ints((7, 600004)), (0, ))
red('cols few rows ragged [2, 5, 400001]', ints((2, 5, 400001)), (1, ))
red('cols narrow [2^21+3, 3]', ints(((1 << 21) + 3, 3)), (0, ))
red('cols narrow words [2^21, 6] i16', ints((1 << 21, 6), torch.int16), (0, ))
red('cols narrow words [2^20, 8] u8', ints((1 << 20, 8), torch.uint8, 0, 3), (0, ))
red('cols mid axis [64, 512, 256]', ints((64, 512, 256)), (1, ))
# multi-pass plans and the generic kernel
red('cols small + inner first [16, 16, 64, 64, 4] (1, 3)', ints((16, 16, 64, 64, 4)), (1, 3))
red('inner first [32, 64, 32, 128] (0, 2)', ints((32, 64, 32, 128)), (0, 2))
red('outer first [4096, 64, 8] (0, 2)', ints((4096, 64, 8)), (0, 2))
red('two pass [6, 50, 4096] (0, 2)', ints((6, 50, 4096)), (0, 2))
red('generic stepped view', ints((300, 70, 50))[::2, :, ::3], (0, 2))
# arg reductions
for name, shape, axis in (('arg rows', (5000, 256), 1), ('arg long rows', (3, 1 << 20), 1), ('arg cols', (5000, 256), 0),
                          ('arg few rows', (8, 200000), 0), ('arg warp per long row', (5000, 9000), 1)):
    x = torch.rand(shape, device=dev, generator=g)
    assert torch.equal(db.argmax(x, axis=axis).long(), x.argmax(dim=axis)), name
    done.append(name)
# row max + row sum
for name, shape in (('max_sum group', (512, 8192)), ('max_sum tiny', (140000, 33)), ('max_sum tiny padded', (70001, 64)),
                    ('max_sum cluster', (8, 1 << 19)), ('max_sum cluster 1', (700, 33000))):
    x = ints(shape).float()
    if name == 'max_sum tiny padded':
        x = torch.cat([x.reshape(-1)[:1], x.reshape(-1)])[1:].view(shape)      # base off by one element
    mx, sm = db.rows_max_sum(x)
    assert torch.equal(mx, x.amax(1)) and torch.equal(sm, x.sum(1)), name
    done.append(name)
# scatters
n = (1 << 22) + 77
for name, nbins, op, weighted in (('scatter smem sum', 256, ADD, False), ('scatter smem max values', 5000, MAX, True),
                                  ('scatter smem min values', 256, MIN, True), ('scatter global values', 1 << 16, ADD, True),
                                  ('scatter partition 2^20', 1 << 20, ADD, False), ('scatter partition 2^16', 1 << 16, ADD, False),
                                  ('scatter weighted large', 1 << 20, ADD, True)):
    keys = ints((n, ), torch.int32, -2, nbins + 2)
    w = ints((n, ), torch.int32, 0, 5) if weighted else None
    bins = torch.zeros(nbins, dtype=torch.int32, device=dev)
    db.wcr_scatter(op, bins, keys, w)
    ok = (keys >= 0) & (keys < nbins)
    ref = torch.zeros(nbins, dtype=torch.int32, device=dev)
    vals = w[ok] if weighted else torch.ones(int(ok.sum()), dtype=torch.int32, device=dev)
    ref.scatter_reduce_(0, keys[ok].long(), vals, {ADD: 'sum', MAX: 'amax', MIN: 'amin'}[op], include_self=True)
    assert torch.equal(bins, ref), name
    done.append(name)
for name, keys in (('scatter hot-key cache (all equal)', torch.full((n, ), 77, dtype=torch.int32, device=dev)),
                   ('scatter sorted keys (gated global)', torch.sort(ints((n, ), torch.int32, 0, 1 << 20)).values)):
    bins = torch.zeros(1 << 20, dtype=torch.int32, device=dev)
    db.wcr_scatter(ADD, bins, keys)
    assert torch.equal(bins, torch.bincount(keys, minlength=1 << 20).int()), name
    done.append(name)
# device headers through the native harnesses
h = ctypes.CDLL(os.path.join(ROOT, 'tests', 'native', '_built', 'libfused_reduce_harness.so'))
h.fused_reduce_harness_run.restype = ctypes.c_int
h.fused_reduce_harness_run.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong,
                                       ctypes.c_ulonglong, ctypes.c_ulonglong, ctypes.c_int, ctypes.c_void_p,
                                       ctypes.c_ulonglong, ctypes.c_void_p]
ws = db.Workspace(dev)
st = torch.cuda.current_stream().cuda_stream
for case, rows, cols in ((42, 600, 5000), (42, 3, 100000), (42, 40000, 64), (3, 600, 5001), (52, 300, 5000), (52, 64, 70000), (13, 4097, 41)):
    a = ints((rows, cols), torch.int64)
    w = ints((cols, ), torch.int64)
    rowwise = case in (42, 3)
    out = torch.zeros(rows if rowwise else cols, dtype=torch.int64, device=dev)
    rc = h.fused_reduce_harness_run(case, a.data_ptr(), w.data_ptr(), out.data_ptr(), 1, rows, cols, 1, ws.ptr, ws.nbytes, st)
    assert rc == 0
    assert torch.equal(out, (a * w[None, :]).sum(1 if rowwise else 0) + 11), ('fused', case, rows, cols)
    done.append('fused harness case %d [%d, %d]' % (case, rows, cols))
torch.cuda.synchronize()
print('all_kernels_check.py: %d calls ok' % len(done))
for d in done:
    print('  ' + d)
