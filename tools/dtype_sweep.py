"""Sweep of element types over the main shapes (full / rows / columns, 1 GiB inputs) to find per-type cliffs:
register counts differ per instantiation, and so does what fits an SM."""
import json, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dace_b200 as db
try:
    PEAK = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']
except Exception:
    PEAK = 6500.0
dev = torch.device('cuda:0')
OPS = {'sum': ('lambda a, b: a + b', 0), 'max': ('lambda a, b: max(a, b)', None)}
BYTES = 1 << 30
def bench(fn, reps=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
dtypes = [torch.float32, torch.float64, torch.int32, torch.int64, torch.bfloat16, torch.float16, torch.uint8, torch.int16]
only = sys.argv[1:]
print('%-10s %-22s %-4s %9s %8s %6s' % ('dtype', 'shape', 'op', 'ms', 'GB/s', 'frac'))
for dt in dtypes:
    name = str(dt).replace('torch.', '')
    if only and name not in only:
        continue
    es = torch.empty((), dtype=dt).element_size()
    n = BYTES // es
    shapes = [((n, ), None), ((n // 4096, 4096), (1, )), ((n // 256, 256), (1, )), ((n // 4096, 4096), (0, )),
              ((n // 128, 128), (0, )), ((64, n // (64 * 1024), 1024), (1, )), ((256, 64, n // (256 * 64)), (0, 2))]
    for shape, axes in shapes:
        if dt.is_floating_point:
            x = torch.rand(shape, device=dev, dtype=torch.float32).to(dt) if es < 4 else torch.rand(shape, device=dev, dtype=dt)
        else:
            x = torch.randint(0, 100, shape, device=dev, dtype=dt)
        for op, (lam, ident) in OPS.items():
            if ident is None:
                ident = -3e38 if dt.is_floating_point and es >= 4 else (-60000 if dt.is_floating_point else 0)
            try:
                out = db.reduce(lam, x, axis=axes, identity=ident)
                ms = bench(lambda: db.reduce(lam, x, out, axis=axes, identity=ident))
            except Exception as ex:
                print('%-10s %-22s %-4s %s' % (name, list(shape), op, type(ex).__name__ + ': ' + str(ex)[:60]), flush=True)
                continue
            gbs = x.numel() * es / ms / 1e6
            print('%-10s %-22s %-4s %9.3f %8.0f %6.2f%s' % (name, '%s %s' % (list(shape), axes or ''), op, ms, gbs, gbs / PEAK,
                                                          '   <--' if gbs / PEAK < 0.85 else ''), flush=True)
        del x
