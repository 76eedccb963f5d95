"""Ad-hoc probe: transposed / permuted views, with and without the memory-order axis sort."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
ADD = 'lambda a, b: a + b'
cases = [('A.T sum axis=1  [8192,8192]', lambda: torch.rand(8192, 8192, device=dev).t(), (1, )),
         ('A.T sum axis=0  [8192,8192]', lambda: torch.rand(8192, 8192, device=dev).t(), (0, )),
         ('NHWC->NCHW view sum (0,2,3) [64,256,56,56]', lambda: torch.rand(64, 56, 56, 256, device=dev).permute(0, 3, 1, 2), (0, 2, 3)),
         ('NHWC->NCHW view sum (1,)    [64,256,56,56]', lambda: torch.rand(64, 56, 56, 256, device=dev).permute(0, 3, 1, 2), (1, ))]
for name, mk, axes in cases:
    x = mk()
    nbytes = x.numel() * x.element_size()
    ref = x.double().sum(axes)
    for mode in ('sorted', 'unsorted'):
        if mode == 'unsorted': os.environ['DACE_B200_NO_AXIS_SORT'] = '1'
        else: os.environ.pop('DACE_B200_NO_AXIS_SORT', None)
        out = db.reduce(ADD, x, axis=axes, identity=0)
        torch.cuda.synchronize()
        reps = 10 if mode == 'sorted' else 3
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): db.reduce(ADD, x, out, axis=axes, identity=0)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print('%-46s %-9s %8.3f ms %8.0f GB/s %s' % (name, mode, ms, nbytes / ms / 1e6, 'ok' if torch.allclose(out.double(), ref, rtol=1e-4) else 'BAD'), flush=True)
    del x
