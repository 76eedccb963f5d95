"""Ad-hoc sweep: column-sum device time, wide-row kernel vs cluster kernel, over a grid of shapes."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
fb = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
add = 'lambda a, b: a + b'
def t(fn, k=12):
    for _ in range(2): fn()
    tot = 0.0
    for _ in range(k):
        fb.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / k * 1e3
print('rows cols MiB wide_us old_us row_sum_us')
for rows in (256, 1024, 4096, 16384, 65536, 1 << 19):
    for cols in (1024, 4096, 16384):
        mib = rows * cols * 4 >> 20
        if mib < 16 or mib > 8192: continue
        a = torch.rand(rows, cols, device=dev); out = torch.zeros(cols, device=dev); out_r = torch.zeros(rows, device=dev)
        os.environ.pop('DACE_B200_NO_WIDE_COLS', None)
        ws = []
        for wc in (4, 8, 16, 32):
            os.environ['DACE_B200_WIDE_WC'] = str(wc)
            ws.append(t(lambda: db.reduce(add, a, out, axis=0, identity=0)))
        w = min(ws)
        os.environ['DACE_B200_NO_WIDE_COLS'] = '1'
        o = t(lambda: db.reduce(add, a, out, axis=0, identity=0))
        r = t(lambda: db.reduce(add, a, out_r, axis=1, identity=0))
        print('%7d %6d %5d  wc4/8/16/32: %s  old %7.1f  row %7.1f' % (rows, cols, mib, ' '.join('%7.1f' % x for x in ws), o, r), flush=True)
        del a
