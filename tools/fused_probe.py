"""Ad-hoc probe: s[i] = sum_j exp(A[i, j] - mx[i]) as a DaCe program, fused (one generated kernel with a load
functor) vs unfused (the reference code generator's map kernel writing a transient + 'B200' Reduce) vs the
reference's own MapReduceFusion of the same program (oracle/_ref, WCR map with atomics).
Device-resident arguments; wall clock around the program call with the stream synchronised (the generated
program synchronises at its end)."""
import os, sys, time, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader            # loader of DaCe build products (test infrastructure)
built = os.path.join(ROOT, 'tests', 'dace_e2e', '_built')
dev = torch.device('cuda:0')
for rows, n in ((65536, 8192), (8192, 65536), (1 << 20, 512)):
    a = torch.randn(rows, n, device=dev)
    mx = a.max(dim=1).values.contiguous()
    s = torch.zeros(rows, device=dev)
    ref = None
    for name in ('b200_fused_softmax_den', 'b200_unfused_softmax_den', 'refgpu_softmax_den_mapreducefusion'):
        if name.startswith('refgpu') and rows > 65535:
            continue                      # the reference's map kernel cannot launch: grid.y = rows
        p = ref_loader.RefProgram(name, built if name.startswith('b200') else ref_loader.REF_DIR)
        for _ in range(2):
            p(A=a.data_ptr(), mx=mx.data_ptr(), s=s.data_ptr(), N=rows, M=n)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = 5
        for _ in range(reps):
            p(A=a.data_ptr(), mx=mx.data_ptr(), s=s.data_ptr(), N=rows, M=n)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) / reps * 1e3
        if ref is None:
            ref = torch.exp(a[:64].double() - mx[:64].double()[:, None]).sum(1)
        ok = torch.allclose(s[:64].double(), ref, rtol=2e-5)
        nbytes = a.numel() * 4
        print('[%d, %d] f32  %-26s %8.3f ms  %7.0f GB/s of input  %s' % (rows, n, name, ms, nbytes / ms / 1e6, 'ok' if ok else 'BAD'), flush=True)
        p.close()
    del a
