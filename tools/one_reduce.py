"""One reduction, a few launches (for ncu): python tools/one_reduce.py DTYPE OP AXES SHAPE...   e.g. float32 sum 1 1048576 256"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dt = getattr(torch, sys.argv[1]); op = sys.argv[2]
axes = None if sys.argv[3] == 'all' else tuple(int(a) for a in sys.argv[3].split(','))
shape = tuple(int(a) for a in sys.argv[4:])
dev = torch.device('cuda:0')
x = torch.rand(shape, device=dev).to(dt) if dt.is_floating_point else torch.randint(0, 100, shape, device=dev, dtype=dt)
lam, ident = ('lambda a, b: a + b', 0) if op == 'sum' else ('lambda a, b: max(a, b)', -60000 if dt.is_floating_point else 0)
out = db.reduce(lam, x, axis=axes, identity=ident)
for _ in range(4): db.reduce(lam, x, out, axis=axes, identity=ident)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): db.reduce(lam, x, out, axis=axes, identity=ident)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print('%s %s %s %s: %.3f ms  %.0f GB/s' % (sys.argv[1], op, axes, list(shape), ms, x.numel() * x.element_size() / ms / 1e6))
