"""Sweep of dace_b200.argmax over shapes / axes / types (1 GiB inputs) to find cliffs."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
BYTES = 1 << 30
print('%-9s %-24s %-5s %9s %8s' % ('dtype', 'shape', 'axis', 'ms', 'GB/s'))
for dt in (torch.float32, torch.float64, torch.int32, torch.int64):
    es = torch.empty((), dtype=dt).element_size()
    n = BYTES // es
    for shape, axis in (((1, n), 1), ((n // 4096, 4096), 1), ((n // 256, 256), 1), ((n // 64, 64), 1), ((n // 4096, 4096), 0),
                        ((n // 128, 128), 0), ((8, n // 8), 0), ((64, n // (64 * 1024), 1024), 1)):
        x = torch.rand(shape, device=dev, dtype=dt) if dt.is_floating_point else torch.randint(0, 1 << 30, shape, device=dev, dtype=dt)
        try:
            idx = db.argmax(x, axis=axis)
            ok = torch.equal(idx.long(), x.argmax(dim=axis))
            for _ in range(2): db.argmax(x, axis=axis)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5): db.argmax(x, axis=axis)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            gbs = x.numel() * es / ms / 1e6
            print('%-9s %-24s %-5d %9.3f %8.0f %s%s' % (str(dt).replace('torch.', ''), list(shape), axis, ms, gbs, '' if ok else 'MISMATCH',
                                                       '   <--' if gbs < 5400 else ''), flush=True)
        except Exception as ex:
            print('%-9s %-24s %-5d %s' % (str(dt).replace('torch.', ''), list(shape), axis, str(ex)[:80]), flush=True)
        del x
