"""Sweep of dace_b200.rows_max_sum (row max + row sum in one pass) over row lengths and types, 1 GiB inputs."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
BYTES = 1 << 30
print('%-9s %-22s %9s %8s' % ('dtype', 'shape', 'ms', 'GB/s'))
for dt in (torch.bfloat16, torch.float32, torch.float16):
    es = torch.empty((), dtype=dt).element_size()
    n = BYTES // es
    for cols in (8192, 4096, 1024, 256, 100, 64, 33, 1 << 20):
        rows = n // cols
        x = torch.rand(rows, cols, device=dev).to(dt)
        try:
            mx, sm = db.rows_max_sum(x)
            ok = torch.equal(mx.float(), x.float().amax(1)) and torch.allclose(sm.double(), x.double().sum(1), rtol=1e-4)
            for _ in range(2): db.rows_max_sum(x, mx, sm)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5): db.rows_max_sum(x, mx, sm)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            gbs = x.numel() * es / ms / 1e6
            print('%-9s %-22s %9.3f %8.0f %s%s' % (str(dt).replace('torch.', ''), [rows, cols], ms, gbs, '' if ok else 'MISMATCH', '   <--' if gbs < 5400 else ''), flush=True)
        except Exception as ex:
            print('%-9s %-22s %s' % (str(dt).replace('torch.', ''), [rows, cols], str(ex)[:90]), flush=True)
        del x
