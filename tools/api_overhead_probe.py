"""Ad-hoc probe: host-side cost of one dace_b200.reduce call (launch only, tiny input, no sync in the loop)."""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
x = torch.rand(64, 64, device='cuda'); out = torch.zeros(64, device='cuda')
ADD = 'lambda a, b: a + b'
for _ in range(100): db.reduce(ADD, x, out, axis=1, identity=0)
torch.cuda.synchronize()
for n in (2000, 2000):
    t = time.perf_counter()
    for _ in range(n): db.reduce(ADD, x, out, axis=1, identity=0)
    dt = time.perf_counter() - t
    torch.cuda.synchronize()
    print('reduce: %.2f us per call (host side)' % (dt / n * 1e6))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(2000): db.reduce(ADD, x, out, axis=1, identity=0)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats('tottime').print_stats(12)
