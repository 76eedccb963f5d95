import sys, torch, itertools
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
base = torch.rand(1024, 512, 512, device='cuda')
for name, x in (('window [:, :500, :500]', base[:, :500, :500]), ('2-D window', base.view(1024*512, 512)[:, :500]), ('4-D interior', base.view(32,32,512,512)[:, :, 1:-1, 1:-1])):
    for axis in range(x.dim()):
        try:
            idx = db.argmax(x, axis=axis)
            ok = torch.equal(idx.long(), x.argmax(dim=axis))
            db.argmax(x, axis=axis); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3): db.argmax(x, axis=axis)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 3
            print('%-26s axis %d  %8.3f ms %7.0f GB/s %s' % (name, axis, ms, x.numel() * 4 / ms / 1e6, '' if ok else 'MISMATCH'), flush=True)
        except Exception as ex:
            print('%-26s axis %d  %s' % (name, axis, str(ex)[:100]), flush=True)
