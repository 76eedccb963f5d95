"""Layouts that end in the generic strided kernel (steps, flips, padded pitches): GB/s of the bytes the kernel must
touch (useful elements x element size; sector efficiency of a step-2 view is 50 % by construction)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
ADD = 'lambda a, b: a + b'
base = torch.rand(16384, 16384, device=dev)           # 1 GiB
cases = [
    ('rows of a padded pitch [16384, :16000]', base[:, :16000], (1, )),
    ('cols of a padded pitch [16384, :16000]', base[:, :16000], (0, )),
    ('full of a padded pitch', base[:, :16000], None),
    ('every 2nd row, rows', base[::2, :], (1, )),
    ('every 2nd row, cols', base[::2, :], (0, )),
    ('every 2nd column, rows', base[:, ::2], (1, )),
    ('every 2nd column, cols', base[:, ::2], (0, )),
    ('flipped rows (negative stride), rows', base.flip(0) if False else base[torch.arange(16383, -1, -1, device=dev)] , (1, )),
    ('transposed view, rows (= cols of the base)', base.t(), (1, )),
    ('transposed view, cols (= rows of the base)', base.t(), (0, )),
    ('3-D permuted [128, 512, 4096] -> (2, 0, 1) reduce axis 0', base.view(128, 512, 4096).permute(2, 0, 1), (0, )),
    ('3-D permuted reduce axes (0, 2)', base.view(128, 512, 4096).permute(1, 0, 2), (0, 2)),
    ('diagonal-ish: [::3, 5::7]', base[::3, 5::7], (1, )),
]
print('%-62s %9s %9s' % ('case', 'ms', 'GB/s'))
for name, x, axes in cases:
    try:
        out = db.reduce(ADD, x, axis=axes, identity=0)
        ref = x.double().sum(axes) if axes is not None else x.double().sum()
        ok = torch.allclose(out.double().reshape(ref.shape), ref, rtol=1e-4)
        for _ in range(2): db.reduce(ADD, x, out, axis=axes, identity=0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): db.reduce(ADD, x, out, axis=axes, identity=0)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        print('%-62s %9.3f %9.0f %s' % (name, ms, x.numel() * 4 / ms / 1e6, '' if ok else 'MISMATCH'), flush=True)
    except Exception as ex:
        print('%-62s %s' % (name, str(ex)[:80]), flush=True)
