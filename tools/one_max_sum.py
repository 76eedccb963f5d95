import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dt = getattr(torch, sys.argv[1]); rows, n = int(sys.argv[2]), int(sys.argv[3])
x = torch.rand(rows, n, device='cuda').to(dt)
mx, sm = db.rows_max_sum(x)
for _ in range(3): db.rows_max_sum(x, mx, sm)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): db.rows_max_sum(x, mx, sm)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print('max_sum %s [%d, %d]: %.3f ms %.0f GB/s' % (sys.argv[1], rows, n, ms, x.numel() * x.element_size() / ms / 1e6))
