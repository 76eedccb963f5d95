"""Sweep of awkward shapes (tiny rows, few columns, odd sizes, unaligned pointers, interleaved axes) to find
performance cliffs; ~1 GiB f32 inputs, back-to-back launches, CUDA events."""
import json, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dace_b200 as db
try:
    PEAK = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']
except Exception:
    PEAK = 6500.0
dev = torch.device('cuda:0')
ADD = 'lambda a, b: a + b'
def bench(fn, reps=5, graph=False):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if graph:
        # small inputs: the Python call (~35 us) is longer than the kernels, so time a CUDA graph of `reps` calls
        # (the library allocates nothing and never synchronises, so its calls can be captured)
        reps = 20
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2): fn()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for _ in range(reps): fn()
        torch.cuda.synchronize()
        g.replay(); torch.cuda.synchronize()
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
N = 1 << 28
cases = [
    ('rows n=33', (N // 33, 33), (1, )), ('rows n=100', (N // 100, 100), (1, )), ('rows n=1000', (N // 1000, 1000), (1, )),
    ('rows n=4097', (N // 4097, 4097), (1, )), ('rows 3 x huge', (3, N // 3), (1, )), ('rows 200 x 1.3M', (200, N // 200), (1, )),
    ('cols K=3', (N // 3, 3), (0, )), ('cols K=17', (N // 17, 17), (0, )), ('cols K=100', (N // 100, 100), (0, )),
    ('cols K=1000', (N // 1000, 1000), (0, )), ('cols K=65536', (N // 65536, 65536), (0, )), ('cols 7 rows', (7, N // 7), (0, )),
    ('full 2^28+1', (N + 1, ), None), ('full 2^28-3', (N - 3, ), None),
    ('[64,256,R2=16] (0,2)', (64, 256, N // (64 * 256)), (0, 2)), ('[4096,512,R2=8] (0,2)', (4096, 512, 8), (0, 2)),
    ('[4096,8,R2=512] (0,2)', (4096, 8, 512), (0, 2)), ('[32,64,32,1024] (0,2)', (32, 64, 32, 1024), (0, 2)),
    ('[32,64,32,1024] (1,3)', (32, 64, 32, 1024), (1, 3)), ('[64,64,64,64,4] (1,3)', (64, 64, 64, 64, 4), (1, 3)),
    ('[N,M,K] mid axis 1', (1024, 256, 1024), (1, )),
]
ONLY = sys.argv[1:]       # case names (substring match): run just those, e.g. for an ncu capture
if ONLY:
    cases = [c for c in cases if any(o in c[0] for o in ONLY)]
print('%-28s %10s %9s %6s' % ('case', 'ms', 'GB/s', 'frac'))
for name, shape, axes in cases:
    x = torch.rand(shape, device=dev)
    out = db.reduce(ADD, x, axis=axes, identity=0)
    small = x.numel() * 4 <= (256 << 20)
    ms = bench(lambda: db.reduce(ADD, x, out, axis=axes, identity=0), graph=small)
    gbs = x.numel() * 4 / ms / 1e6
    ok = torch.allclose(out.double(), x.double().sum(axes) if axes is not None else x.double().sum(), rtol=1e-4)
    print('%-28s %10.3f %9.0f %6.2f %s%s' % (name, ms, gbs, gbs / PEAK, '' if ok else 'BAD', ' (graph of 20)' if small else ''), flush=True)
    del x, out
if ONLY:
    sys.exit(0)
# unaligned pointers: a view that starts 1 element into the allocation
base = torch.rand(N + 8, device=dev)
for off in (1, 3):
    v = base[off:off + N]
    out = torch.zeros((), device=dev)
    ms = bench(lambda: db.reduce(ADD, v, out, axis=None, identity=0))
    print('%-28s %10.3f %9.0f %6.2f' % ('full, pointer + %d elem' % off, ms, N * 4 / ms / 1e6, N * 4 / ms / 1e6 / PEAK), flush=True)
m = base[1:1 + (N // 4096) * 4096].view(N // 4096, 4096)
outr = torch.zeros(N // 4096, device=dev)
ms = bench(lambda: db.reduce(ADD, m, outr, axis=1, identity=0))
print('%-28s %10.3f %9.0f %6.2f' % ('rows n=4096, pointer + 1', ms, m.numel() * 4 / ms / 1e6, m.numel() * 4 / ms / 1e6 / PEAK), flush=True)
