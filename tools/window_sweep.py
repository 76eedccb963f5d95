"""Reductions over WINDOWS of a larger 3-D array (padding on two levels: [1024, 512, 512] f32, window
[:, :500, :500] = 0.95 of the bytes): every axis combination, GB/s of the useful bytes."""
import os, sys, itertools, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
ADD = 'lambda a, b: a + b'
base = torch.rand(1024, 512, 512, device=dev)
views = {'window [:, :500, :500]': base[:, :500, :500], 'window [3:1000, 7:, :500]': base[3:1000, 7:, :500],
         'window [:, :500, :] (one padded level)': base[:, :500, :]}
base4 = base.view(32, 32, 512, 512)
views4 = {'4-D interior [:, :, 1:-1, 1:-1]': base4[:, :, 1:-1, 1:-1]}
print('%-42s %-10s %9s %9s' % ('view', 'axes', 'ms', 'GB/s'))
ONLY4 = '--4d' in sys.argv or '--3level' in sys.argv
if '--3level' in sys.argv:   # a pitch after every row, plane and cube: the outermost axis is peeled
    b4 = base.view(16, 64, 64, 4096)
    views4 = {'3 padded levels [:, :60, :60, :4000]': b4[:, :60, :60, :4000],      # 57.6 MB slices
              '3 padded levels [:, :60, :60, :1000]': b4[:, :60, :60, :1000]}      # 14.4 MB slices
ALL = [(vn, x, axes) for vn, x in ({} if ONLY4 else views).items() for r in (1, 2, 3) for axes in itertools.combinations(range(3), r)]
ALL += [(vn, x, axes) for vn, x in views4.items() for axes in ((2, 3), (0, 2, 3), (3, ), (1, 2, 3), (0, 1, 2, 3), (0, 1), (1, ), (0, ), (2, ), (0, 2), (1, 3), (0, 3), (1, 2), (0, 1, 3), (0, 1, 2))]
for vname, x, axes in ALL:
    if True:
        if True:
            try:
                out = db.reduce(ADD, x, axis=axes, identity=0)
                ref = x.double().sum(axes)
                ok = torch.allclose(out.double().reshape(ref.shape), ref, rtol=1e-4)
                db.reduce(ADD, x, out, axis=axes, identity=0)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(3): db.reduce(ADD, x, out, axis=axes, identity=0)
                e1.record(); torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 3
                gbs = x.numel() * 4 / ms / 1e6
                print('%-42s %-10s %9.3f %9.0f %s%s' % (vname, axes, ms, gbs, '' if ok else 'MISMATCH', '   <--' if gbs < 3000 else ''), flush=True)
            except Exception as ex:
                print('%-42s %-10s %s' % (vname, axes, str(ex)[:80]), flush=True)
