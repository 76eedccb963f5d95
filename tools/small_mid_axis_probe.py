"""Middle axis with very few elements ([N, 3, M] over axis 1: channel folds) and other short reduced axes, 1 GiB f32."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
ADD = 'lambda a, b: a + b'
n = 1 << 28
print('%-34s %-8s %9s %9s' % ('shape', 'axes', 'ms', 'GB/s'))
for shape, axes in (((4096, 4, 16384), (1, )), ((5461, 3, 16384), (1, )), ((1 << 20, 4, 64), (1, )), ((64, 2, 1 << 21), (1, )),
                    ((1 << 16, 8, 512), (1, )), ((1 << 22, 2, 32), (1, )), ((3, 1 << 20, 85), (0, )), ((2, 1 << 27), (0, )),
                    ((1 << 26, 4), (1, )), ((1 << 27, 2), (1, )), ((4, 1 << 13, 1 << 13), (0, 2)), ((1 << 13, 4, 1 << 13), (0, 2))):
    x = torch.rand(shape, device=dev)
    try:
        out = db.reduce(ADD, x, axis=axes, identity=0)
        ok = torch.allclose(out.double(), x.double().sum(axes), rtol=1e-4)
        db.reduce(ADD, x, out, axis=axes, identity=0); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): db.reduce(ADD, x, out, axis=axes, identity=0)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        gbs = (x.numel() + out.numel()) * 4 / ms / 1e6
        print('%-34s %-8s %9.3f %9.0f %s%s' % (list(shape), axes, ms, gbs, '' if ok else 'MISMATCH', '   <--' if gbs < 4000 else ''), flush=True)
    except Exception as ex:
        print('%-34s %-8s %s' % (list(shape), axes, str(ex)[:80]), flush=True)
    del x
