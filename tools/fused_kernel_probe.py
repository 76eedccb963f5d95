"""Ad-hoc probe: dace_b200::fused_rows / fused_cols (include/dace_b200_fused_reduce.cuh) with the exp(x - m[r])
functor of the native test harness, CUDA events, inputs larger than L2."""
import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dace_b200 as db
h = ctypes.CDLL(os.path.join(ROOT, 'tests', 'native', '_built', os.environ.get('FUSED_HARNESS', 'libfused_reduce_harness.so')))
h.fused_reduce_harness_run.restype = ctypes.c_int
h.fused_reduce_harness_run.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong,
                                       ctypes.c_ulonglong, ctypes.c_ulonglong, ctypes.c_int, ctypes.c_void_p,
                                       ctypes.c_ulonglong, ctypes.c_void_p]
dev = torch.device('cuda:0')
ws = db.Workspace(dev)
s = torch.cuda.current_stream().cuda_stream
CASES = ((0, 'fused_rows'), (40, 'fused_rows v4'), (41, 'fused_rows v8'), (10, 'fused_cols'), (50, 'fused_cols v4'), (51, 'fused_cols v8'))
SHAPES = ((65536, 8192), (8192, 65536), (1 << 20, 512), (1 << 22, 128), (4, 1 << 27), (1 << 19, 1024))
if len(sys.argv) == 4:      # one case, one shape (for ncu): python tools/fused_kernel_probe.py CASE ROWS N
    CASES = tuple(c for c in CASES if c[0] == int(sys.argv[1]))
    SHAPES = ((int(sys.argv[2]), int(sys.argv[3])), )
for case, name in CASES:
    for rows, n in SHAPES:
        x = torch.randn(rows, n, device=dev)
        m = torch.zeros(rows, device=dev)
        out = torch.zeros(rows if case in (0, 40, 41) else n, device=dev)
        def run():
            rc = h.fused_reduce_harness_run(case, x.data_ptr(), m.data_ptr(), out.data_ptr(), 1, rows, n, 1, ws.ptr, ws.nbytes, s)
            assert rc == 0
        for _ in range(3): run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): run()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print('%-13s [%8d, %9d] f32 exp(x - m[r])  %7.3f ms  %6.0f GB/s' % (name, rows, n, ms, x.numel() * 4 / ms / 1e6), flush=True)
        del x
