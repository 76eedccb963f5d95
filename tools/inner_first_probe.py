"""Single-shape probe for ncu: [64, 64, 64, 64, 4] f32 over axes (1, 3) - inner-first plan = cols_small_kernel over
262144 blocks of [64, 4], then the column kernels on the intermediates."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
x = torch.rand(64, 64, 64, 64, 4, device=dev)
out = torch.zeros(64, 64, 4, device=dev)
for _ in range(5):
    db.reduce('lambda a, b: a + b', x, out, axis=(1, 3), identity=0)
torch.cuda.synchronize()
assert torch.allclose(out.double(), x.double().sum((1, 3)), rtol=1e-4)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    db.reduce('lambda a, b: a + b', x, out, axis=(1, 3), identity=0)
e1.record(); torch.cuda.synchronize()
print('%.4f ms per call, %.0f GB/s' % (e0.elapsed_time(e1) / 10, x.numel() * 4 / (e0.elapsed_time(e1) / 10) / 1e6))
