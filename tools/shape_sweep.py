"""Sweep: every kernel family x dtype on ~1 GiB inputs (far larger than L2), back-to-back launches, CUDA events.
Prints achieved GB/s of input and the fraction of the measured HBM copy peak (MEASURED_PEAKS.json)."""
import json, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dace_b200 as db
try:
    PEAK = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']
except Exception:
    PEAK = 6500.0
dev = torch.device('cuda:0')
ADD, MAX = 'lambda a, b: a + b', 'lambda a, b: max(a, b)'
DT = {'f32': torch.float32, 'f64': torch.float64, 'i32': torch.int32, 'i64': torch.int64, 'bf16': torch.bfloat16,
      'f16': torch.float16, 'u8': torch.uint8, 'i16': torch.int16}
def make(shape, dt):
    if dt.is_floating_point:
        return torch.rand(shape, device=dev, dtype=torch.float32).to(dt)
    return torch.randint(0, 100, shape, device=dev, dtype=dt)
def bench(fn, reps=10):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
GIB = 1 << 30
print('%-34s %-5s %-4s %9s %9s %6s' % ('layout', 'dtype', 'op', 'ms', 'GB/s', 'frac'))
for name, dt in DT.items():
    es = torch.empty(0, dtype=dt).element_size()
    n = GIB // es
    layouts = [('full [%d]' % n, (n, ), None),
               ('rows [65536, %d] axis 1' % (n // 65536), (65536, n // 65536), (1, )),
               ('cols [%d, 4096] axis 0' % (n // 4096), (n // 4096, 4096), (0, )),
               ('two-pass [64, 256, %d] axes (0,2)' % (n // 16384), (64, 256, n // 16384), (0, 2))]
    for lname, shape, axes in layouts:
        x = make(shape, dt)
        for opname, lam, ident in (('sum', ADD, 0), ('max', MAX, -3.0e38 if dt.is_floating_point else -2**7 if es == 1 else -2**15)):
            if opname == 'max' and lname.startswith('two-pass') and name not in ('f32', 'i32'):
                continue
            try:
                out = db.reduce(lam, x, axis=axes, identity=ident if dt.is_floating_point or opname == 'sum' else 0)
                ms = bench(lambda: db.reduce(lam, x, out, axis=axes, identity=ident if dt.is_floating_point or opname == 'sum' else 0))
                gbs = x.numel() * es / ms / 1e6
                print('%-34s %-5s %-4s %9.3f %9.0f %6.2f' % (lname, name, opname, ms, gbs, gbs / PEAK), flush=True)
            except Exception as ex:
                print('%-34s %-5s %-4s  unsupported (%s)' % (lname, name, opname, str(ex)[:50]), flush=True)
        del x
