"""Histogram throughput by KEY type (int32 / int64 / uint8 keys), constant increment, 2^26 uniform keys."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
n = 1 << 26
g = torch.Generator(device=dev); g.manual_seed(0)
ADD = 'lambda a, b: a + b'
print('%-7s %-9s %9s %10s' % ('keys', 'nbins', 'ms', 'Gkeys/s'))
for kdt in (torch.int32, torch.int64, torch.uint8):
    for nbins in (256, 1 << 16, 1 << 20):
        if kdt == torch.uint8 and nbins != 256:
            continue
        keys = torch.randint(0, nbins, (n, ), device=dev, dtype=kdt, generator=g)
        bins = torch.zeros(nbins, dtype=torch.int32, device=dev)
        try:
            db.wcr_scatter(ADD, bins, keys)
            ok = torch.equal(bins, torch.bincount(keys.long(), minlength=nbins).int())
            bins.zero_()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3): db.wcr_scatter(ADD, bins, keys)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 3
            print('%-7s %-9d %9.3f %10.1f %s' % (str(kdt).replace('torch.', ''), nbins, ms, n / ms / 1e6, '' if ok else 'MISMATCH'), flush=True)
        except Exception as ex:
            print('%-7s %-9d %s' % (str(kdt).replace('torch.', ''), nbins, str(ex)[:90]), flush=True)
