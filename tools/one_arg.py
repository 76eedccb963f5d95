"""One arg-reduction over the last axis, timed: python tools/one_arg.py DTYPE ROWS N"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dt = getattr(torch, sys.argv[1]); rows, n = int(sys.argv[2]), int(sys.argv[3])
dev = torch.device('cuda:0')
x = torch.rand(rows, n, device=dev).to(dt) if dt.is_floating_point else torch.randint(0, 1000, (rows, n), device=dev, dtype=dt)
idx = db.argmax(x, axis=1)
assert torch.equal(idx.long(), x.argmax(dim=1)) or not dt.is_floating_point
for _ in range(3): db.argmax(x, axis=1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): db.argmax(x, axis=1)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print('argmax %s [%d, %d]: %.3f ms  %.0f GB/s' % (sys.argv[1], rows, n, ms, x.numel() * x.element_size() / ms / 1e6))
