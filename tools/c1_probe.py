"""Ad-hoc probe: C1 (4096x4096 f32 row sum) kernel time under different L2 states."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
a = torch.rand(4096, 4096, device=dev)
out = torch.zeros(4096, device=dev)
fb = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
rb = torch.empty(64 << 20, dtype=torch.float32, device=dev)   # 256 MiB
add = 'lambda a, b: a + b'
def run(): db.reduce(add, a, out, axis=1, identity=0)
for _ in range(5): run()
def t(pre, k=30):
    tot = 0.0
    for _ in range(k):
        pre()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / k * 1e3
print('write flush            %.2f us' % t(lambda: fb.fill_(1)))
print('write + read flush     %.2f us' % t(lambda: (fb.fill_(1), rb.sum())))
print('read-only flush        %.2f us' % t(lambda: rb.sum()))
print('no flush (L2 hot)      %.2f us' % t(lambda: None))
import time
print('write flush + 1ms idle %.2f us' % t(lambda: (fb.fill_(1), torch.cuda.synchronize(), time.sleep(0.001))))
