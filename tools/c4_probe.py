"""Ad-hoc probe: partitioned histogram time per key for L2-resident vs HBM-resident key arrays."""
import sys, torch
import dace_b200 as db
dev = torch.device('cuda:0')
nb = 1 << 20
for lg in (23, 24, 26, 28):
    n = 1 << lg
    keys = torch.randint(0, nb, (n,), device=dev, dtype=torch.int32)
    bins = torch.zeros(nb, device=dev, dtype=torch.int32)
    for _ in range(3):
        db.wcr_scatter('lambda a, b: a + b', bins, keys, None)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    e0.record()
    for _ in range(reps):
        db.wcr_scatter('lambda a, b: a + b', bins, keys, None)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print('n=2^%d  %.4f ms  %.1f Gkeys/s  (scaled to 2^28: %.3f ms)' % (lg, ms, n / ms / 1e6, ms * (1 << 28) / n))
