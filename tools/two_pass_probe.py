"""Ad-hoc probe: [R1, K, R2] reductions (axes (0, 2)), two-pass plan vs the generic strided kernel."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
ADD = 'lambda a, b: a + b'
for shape, dt in (((64, 256, 3136), torch.float32), ((256, 64, 12544), torch.float32), ((32, 1024, 4096), torch.bfloat16),
                  ((128, 2048, 196), torch.float32), ((8, 16, 1 << 22), torch.float32), ((1024, 1024, 256), torch.float32)):
    x = torch.rand(shape, device=dev, dtype=torch.float32).to(dt)
    out = torch.zeros(shape[1], device=dev, dtype=torch.float32)
    nbytes = x.numel() * x.element_size()
    for mode in ('two-pass', 'generic'):
        if mode == 'generic': os.environ['DACE_B200_NO_TWO_PASS'] = '1'
        else: os.environ.pop('DACE_B200_NO_TWO_PASS', None)
        for _ in range(2): db.reduce(ADD, x, out, axis=(0, 2), identity=0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10 if mode == 'two-pass' else 3
        e0.record()
        for _ in range(reps): db.reduce(ADD, x, out, axis=(0, 2), identity=0)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        ok = torch.allclose(out.double(), x.double().sum((0, 2)), rtol=1e-4)
        print('%-22s %-8s %-9s %8.3f ms %8.0f GB/s %s' % (shape, str(dt).split('.')[-1], mode, ms, nbytes / ms / 1e6, 'ok' if ok else 'BAD'), flush=True)
    del x
