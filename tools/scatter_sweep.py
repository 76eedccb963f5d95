"""Sweep of dace_b200.wcr_scatter over operators, value types and bin counts (2^26 uniform int32 keys) to find cliffs."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
n = 1 << 26
g = torch.Generator(device=dev); g.manual_seed(0)
OPS = {'sum': 'lambda a, b: a + b', 'max': 'lambda a, b: max(a, b)', 'min': 'lambda a, b: min(a, b)'}
print('%-5s %-8s %-9s %-7s %9s %10s' % ('op', 'values', 'nbins', 'weights', 'ms', 'Gkeys/s'))
for nbins in (256, 1 << 16, 1 << 20):
    keys = torch.randint(0, nbins, (n, ), device=dev, dtype=torch.int32, generator=g)
    for dt in (torch.int32, torch.float32, torch.int64, torch.float64):
        w = (torch.rand(n, device=dev, generator=g) * 100).to(dt)
        for op, lam in OPS.items():
            for weighted in (False, True):
                if not weighted and op != 'sum':
                    continue
                bins = torch.zeros(nbins, device=dev, dtype=dt)
                vals = w if weighted else None
                try:
                    db.wcr_scatter(lam, bins, keys, vals)
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(3): db.wcr_scatter(lam, bins, keys, vals)
                    e1.record(); torch.cuda.synchronize()
                    ms = e0.elapsed_time(e1) / 3
                    print('%-5s %-8s %-9d %-7s %9.3f %10.1f' % (op, str(dt).replace('torch.', ''), nbins, 'yes' if weighted else 'const', ms, n / ms / 1e6), flush=True)
                except Exception as ex:
                    print('%-5s %-8s %-9d %-7s %s' % (op, str(dt).replace('torch.', ''), nbins, weighted, str(ex)[:70]), flush=True)
        del w
