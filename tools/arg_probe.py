"""Ad-hoc probe: argmax kernel time / bandwidth over the shapes of the benchmark configs."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
fb = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
def t(fn, k=10):
    for _ in range(3): fn()
    tot = 0.0
    for _ in range(k):
        fb.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / k
cases = [('flat 2^30 f32', (1 << 30, ), None, torch.float32), ('rows [65536, 8192] bf16', (65536, 8192), 1, torch.bfloat16),
         ('rows [4096, 4096] f32', (4096, 4096), 1, torch.float32), ('cols [2^19, 4096] f32', (1 << 19, 4096), 0, torch.float32),
         ('cols [65536, 8192] bf16', (65536, 8192), 0, torch.bfloat16), ('mid [64, 4096, 1024] f32', (64, 4096, 1024), 1, torch.float32),
         ('flat 2^28 i32', (1 << 28, ), None, torch.int32), ('rows [2^20, 256] f32', (1 << 20, 256), 1, torch.float32)]
for name, shape, axis, dt in cases:
    a = (torch.rand(shape, device=dev) * 1000).to(dt)
    kept = [s for i, s in enumerate(shape) if axis is not None and i != axis]
    oi = torch.empty(kept, dtype=torch.int32, device=dev); ov = torch.empty(kept, dtype=dt, device=dev)
    ms = t(lambda: db.argmax(a, axis=axis, out_idx=oi, out_val=ov))
    ms_t = t(lambda: torch.argmax(a) if axis is None else torch.argmax(a, dim=axis))
    gb = a.numel() * a.element_size() / 1e9
    print('%-28s %8.3f ms %8.0f GB/s   (torch.argmax %8.3f ms)' % (name, ms, gb / ms * 1e3, ms_t), flush=True)
    del a
