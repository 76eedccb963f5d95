#!/usr/bin/env python
"""Benchmark of the Reduce / WCR hot path on B200 (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c3max|c4|c5|c1] [--impl reference]

One "step" = one pass of the hot path over one batch of synthetic input already
resident in HBM.  Default workload at every N is BASELINE config C2 (full sum of
2^30 float32 to a scalar, per GPU); with N > 1 every rank reduces its own shard
(outer-dimension sharding) and the per-GPU partials are combined by an NCCL
allreduce on the same stream (weak scaling).  One JSON line is printed by rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'reduce_throughput'
UNIT = 'GB/s'
ADD = 'lambda a, b: a + b'
MAX = 'lambda a, b: max(a, b)'

# name -> description, algorithmic bytes per step per GPU (SURVEY.md 8(d) / DESIGN.md)
WORKLOADS = {
    'c2': ('C2 full sum of 2^30 float32 -> scalar', (1 << 30) * 4 + 4),
    'c3': ('C3 fused row max+sum over [65536, 8192] bf16, fp32 accumulate', 65536 * 8192 * 2 + 2 * 65536 * 4),
    'c3max': ('C3 row max over [65536, 8192] bf16 -> fp32 (single Reduce node)', 65536 * 8192 * 2 + 65536 * 4),
    'c4': ('C4 WCR histogram: 2^28 int32 keys -> 2^20 int32 bins (uniform keys)', (1 << 28) * 4 + (1 << 20) * 4),
    'c5': ('C5 column sum of a [2^19, 4096] float32 row shard per GPU', (1 << 19) * 4096 * 4 + 4096 * 4),
    'c1': ('C1 np.sum(A, axis=1) on 4096x4096 float32', 4096 * 4096 * 4 + 4096 * 4),
}


def recorded_traffic(workload):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, 'profiles', 'traffic.json')) as f:
            return json.load(f).get(workload, {}).get('bytes')
    except Exception:
        return None


def peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        with open(path) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md 6.65 TB/s)'


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (NVML poll
    thread every 5 ms; the launch loop releases the GIL inside ctypes calls)."""

    def __init__(self, index=0):
        self.index = index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None

    def _poll(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            bits = {'hw_slowdown': nv.nvmlClocksEventReasonHwSlowdown,
                    'hw_thermal_slowdown': nv.nvmlClocksEventReasonHwThermalSlowdown,
                    'sw_thermal_slowdown': nv.nvmlClocksEventReasonSwThermalSlowdown,
                    'sw_power_cap': nv.nvmlClocksEventReasonSwPowerCap}
            while not self._stop.is_set():
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                for name, bit in bits.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.005)
        except Exception as ex:   # clocks are evidence, not a dependency
            self.error = str(ex)

    def start(self):
        self._thread = threading.Thread(target=self._poll, daemon=True)
        self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)
        sm = sorted(self.samples)
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': self.max_mhz,
                'reasons': sorted(self.reasons), 'samples': len(sm)}


def make_workload(name, dev, rank, torch, db):
    """Returns (step_fn(stream_sync=False) -> result tensor, host_step pieces, description)."""
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    w = {}
    if name == 'c2':
        x = torch.rand(1 << 30, dtype=torch.float32, device=dev, generator=g)
        out = torch.zeros((), dtype=torch.float32, device=dev)
        w['inputs'] = [x]
        w['step'] = lambda: db.reduce(ADD, x, out, axis=None, identity=0)
        w['result'] = out
        w['combine'] = 'sum'
    elif name in ('c3', 'c3max'):
        x = torch.randn(65536, 8192, dtype=torch.float32, device=dev, generator=g).to(torch.bfloat16)
        mx = torch.empty(65536, dtype=torch.float32, device=dev)
        sm = torch.empty(65536, dtype=torch.float32, device=dev)
        w['inputs'] = [x]
        if name == 'c3':
            w['step'] = lambda: db.rows_max_sum(x, mx, sm)
        else:
            w['step'] = lambda: db.reduce(MAX, x, mx, axis=1, identity=-3.0e38)
        w['result'] = mx
        w['combine'] = None   # rows are sharded along the kept axis: no collective
    elif name == 'c4':
        keys = torch.randint(0, 1 << 20, (1 << 28, ), dtype=torch.int32, device=dev, generator=g)
        bins = torch.zeros(1 << 20, dtype=torch.int32, device=dev)
        w['inputs'] = [keys]

        def step():
            bins.zero_()          # part of the step: WCR accumulates into its output
            db.wcr_scatter(ADD, bins, keys)
        w['step'] = step
        w['result'] = bins
        w['combine'] = 'sum'
        w['launches'] = 3   # key sampler + partitioned kernel + gated global-atomic kernel (torch's zero_ not counted)
    elif name == 'c5':
        x = torch.rand(1 << 19, 4096, dtype=torch.float32, device=dev, generator=g)
        out = torch.zeros(4096, dtype=torch.float32, device=dev)
        w['inputs'] = [x]
        w['step'] = lambda: db.reduce(ADD, x, out, axis=0, identity=0)
        w['result'] = out
        w['combine'] = 'sum'
    elif name == 'c1':
        # One 64 MiB matrix would sit in the 126 MB L2 from launch to launch.  Instead of flushing (which
        # forces one event pair per launch: ~5 us of event overhead on a ~14 us kernel) the step walks a
        # ring of 16 distinct matrices (1 GiB): by the time a matrix is read again 960 MiB have gone through L2.
        ring = torch.rand(16, 4096, 4096, dtype=torch.float32, device=dev, generator=g)
        out = torch.zeros(4096, dtype=torch.float32, device=dev)
        pos = [0]
        w['inputs'] = [ring]
        w['e2e_inputs'] = [ring[0]]
        w['e2e_prepare'] = lambda: pos.__setitem__(0, 0)   # the end-to-end step reduces the matrix it uploaded

        def step():
            db.reduce(ADD, ring[pos[0] & 15], out, axis=1, identity=0)
            pos[0] += 1
        w['step'] = step
        w['result'] = out
        w['combine'] = None
        w['l2'] = ('ring of 16 distinct 4096x4096 inputs (1 GiB, 8x the 126 MB L2): consecutive launches read '
                   'different matrices, no flush needed; the K timed launches are replayed from one CUDA graph '
                   '(a ~14 us kernel is otherwise bound by the Python call overhead)')
        w['graph'] = True
    else:
        raise SystemExit('unknown workload %r' % name)
    return w


# ---------------------------------------------------------------------------
# CPU side: the reference's own CPU expansion (oracle/_ref, built from the
# reference by oracle/build_ref.py) on a bounded sample; the C oracle port if
# the reference could not be compiled.  Checker/baseline only.
# ---------------------------------------------------------------------------
def cpu_reference_run(workload, budget_s=12.0, max_reps=5):
    import numpy as np
    from oracle import ref_loader
    cores = os.cpu_count() or 1
    os.environ.setdefault('OMP_NUM_THREADS', str(cores))
    rng = np.random.default_rng(0)
    if workload == 'c2':
        n = 1 << 28
        sample = 'first 2^28 of the 2^30 float32 (1 GiB), reference OpenMP expansion'
        data = {'A': rng.random(n, dtype=np.float32), 'B': np.zeros(1, np.float32), 'D0': n}
        prog, nbytes = 'full_sum_f32_omp', n * 4
    elif workload in ('c3', 'c3max'):
        rows, cols = 8192, 8192
        sample = '[8192, 8192] bf16 slice (128 MiB), reference pure expansion (max, then sum for c3)'
        bits = (rng.standard_normal((rows, cols), dtype=np.float32).view(np.uint32) >> 16).astype(np.uint16)
        data = {'A': bits, 'B': np.zeros(rows, np.float32), 'D0': rows, 'D1': cols}
        prog, nbytes = 'row_max_bf16_f32_pure', rows * cols * 2
    elif workload == 'c4':
        n = 1 << 26
        sample = '2^26 of the 2^28 int32 keys -> 2^20 bins, reference WCR map (omp atomic)'
        data = {'keys': rng.integers(0, 1 << 20, n, dtype=np.int32), 'bins': np.zeros(1 << 20, np.int32), 'N': n}
        prog, nbytes = 'c4_hist_i32', n * 4
    elif workload == 'c5':
        rows, cols = 1 << 16, 4096
        sample = '[2^16, 4096] float32 slice (1 GiB), reference OpenMP expansion'
        data = {'A': rng.random((rows, cols), dtype=np.float32), 'B': np.zeros(cols, np.float32), 'D0': rows, 'D1': cols}
        prog, nbytes = 'col_sum_f32_omp', rows * cols * 4
    else:
        rows, cols = 4096, 4096
        sample = 'full 4096x4096 float32, reference @dace.program np.sum(A, axis=1), OpenMP expansion'
        data = {'A': rng.random((rows, cols), dtype=np.float32), 'B': np.zeros(rows, np.float32), 'N': rows, 'M': cols}
        prog, nbytes = 'c1_npsum_axis1_f32_omp', rows * cols * 4

    if ref_loader.available():
        kind = 'reference'
        p = ref_loader.RefProgram(prog)
        extra = ref_loader.RefProgram('row_sum_bf16_f32_pure') if workload == 'c3' else None

        def run():
            p(**data)
            if extra is not None:
                extra(**data)
    else:
        kind, cores = 'port', 1
        from oracle import oracle as O
        sample += ' [oracle/_ref missing: single-threaded C port]'

        def run():
            if workload == 'c4':
                O.wcr_scatter(O.SUM, data['bins'], data['keys'])
            elif workload in ('c3', 'c3max'):
                O.reduce(O.MAX, data['A'], (1, ), identity=-3e38, in_bf16=True, out_dtype=np.float32)
            elif workload == 'c5':
                O.reduce(O.SUM, data['A'], (0, ), identity=0)
            elif workload == 'c2':
                O.reduce(O.SUM, data['A'], None, identity=0)
            else:
                O.reduce(O.SUM, data['A'], (1, ), identity=0)

    run()  # warm-up (page faults, OpenMP pool)
    times = []
    t_start = time.perf_counter()
    while len(times) < max_reps and (time.perf_counter() - t_start < budget_s or not times):
        t0 = time.perf_counter()
        run()
        times.append(time.perf_counter() - t0)
    best = min(times)
    return {'value': nbytes / best / 1e9, 'unit': UNIT, 'cores': cores, 'kind': kind, 'sample': sample,
            'best_s': best, 'reps': len(times), 'nbytes': nbytes}


# ---------------------------------------------------------------------------
# The reference's own GPU expansions (CUB 'CUDA (device)', 'GPUAuto', generated
# WCR-map kernel) built for sm_100 by oracle/build_ref.py, on device-resident
# arrays of the same shapes: the on-box GPU comparison (--extras only).  The
# generated programs launch on their own streams, so they are timed by wall clock
# with a device synchronise after every call.
# ---------------------------------------------------------------------------
def reference_gpu_run(torch, dev, reps=10):
    from oracle import ref_loader
    if not ref_loader.available():
        return None
    have = ref_loader.manifest()['programs']
    out = {}

    def run(tag, prog, nbytes, make):
        if prog not in have:
            return
        try:
            p = ref_loader.RefProgram(prog)
            keep, kwargs = make()
            for _ in range(2):
                p(**kwargs)
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(reps):
                p(**kwargs)
                torch.cuda.synchronize()
            ms = (time.perf_counter() - t0) / reps * 1e3
            out[tag] = {'GB/s': nbytes / (ms * 1e-3) / 1e9, 'ms': ms, 'program': prog}
            p.close()
            del keep
        except Exception as ex:
            out[tag] = {'error': str(ex)[:200]}
        torch.cuda.empty_cache()

    def c2():
        a = torch.rand(1 << 30, device=dev)
        b = torch.zeros(1, device=dev)
        return (a, b), {'A': a.data_ptr(), 'B': b.data_ptr(), 'D0': a.numel()}

    def c3():
        a = torch.randn(65536, 8192, device=dev, dtype=torch.bfloat16)
        b = torch.zeros(65536, device=dev, dtype=torch.bfloat16)
        return (a, b), {'A': a.data_ptr(), 'B': b.data_ptr(), 'D0': 65536, 'D1': 8192}

    def c5():
        a = torch.rand(1 << 19, 4096, device=dev)
        b = torch.zeros(4096, device=dev)
        return (a, b), {'A': a.data_ptr(), 'B': b.data_ptr(), 'D0': 1 << 19, 'D1': 4096}

    def c4():
        k = torch.randint(0, 1 << 20, (1 << 28, ), device=dev, dtype=torch.int32)
        b = torch.zeros(1 << 20, device=dev, dtype=torch.int32)
        return (k, b), {'keys': k.data_ptr(), 'bins': b.data_ptr(), 'N': k.numel()}

    run('c2_cub', 'refgpu_full_sum_f32_cub', WORKLOADS['c2'][1], c2)
    run('c2_gpuauto', 'refgpu_full_sum_f32_gpuauto', WORKLOADS['c2'][1], c2)
    run('c3_rowsum_bf16_cub', 'refgpu_row_sum_bf16_cub', 65536 * 8192 * 2 + 65536 * 2, c3)
    run('c4_wcr_map_kernel', 'refgpu_hist_i32', WORKLOADS['c4'][1], c4)
    run('c5_shard_colsum_gpuauto', 'refgpu_col_sum_f32_gpuauto', WORKLOADS['c5'][1], c5)
    return out


def run_reference_arm(args):
    """--impl reference: the reference's own CPU expansion of the path on the host cores."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    desc, _ = WORKLOADS[args.workload]
    res = []
    for _ in range(max(1, args.warmup)):
        cpu_reference_run(args.workload, budget_s=0.0, max_reps=1)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res.append(cpu_reference_run(args.workload, budget_s=0.0, max_reps=1))
        if time.perf_counter() - t0 > 150:
            break
    mean_s = sum(r['best_s'] for r in res) / len(res)
    best = res[0]['nbytes'] / mean_s / 1e9     # throughput over exactly the timed steps
    base = dict(res[0])
    base['value'] = best
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': best, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': len(res),
        'warmup': max(1, args.warmup), 'ms_per_step': mean_s * 1e3, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': dtype_of(args.workload), 'data': 'synthetic',
        'config': {'workload': desc, 'sample': base['sample']},
        'cpu_baseline': {k: base[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')},
        'e2e': {'value': best, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line))
    return 0


def dtype_of(workload):
    return {'c2': 'f32', 'c3': 'bf16->f32', 'c3max': 'bf16->f32', 'c4': 'i32', 'c5': 'f32', 'c1': 'f32'}[workload]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=100)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--workload', default='c2', choices=sorted(WORKLOADS))
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--e2e-steps', type=int, default=3)
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    ap.add_argument('--combine', choices=('peer', 'nccl'), default='peer',
                    help='N>1: combine the per-GPU partials with the one-shot NVLink peer-memory kernel (default) or NCCL')
    ap.add_argument('--extras', action='store_true', help='also time the other configs (device-only) into "others"')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'b200' else args.warmup
    if args.impl == 'reference':
        return run_reference_arm(args)

    import torch
    import dace_b200 as db
    from dace_b200 import sharding
    db.load()   # fails loudly if the CUDA library is missing: no fallback
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device; the B200 backend has no CPU fallback')

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', '29531')
        dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)

    desc, bytes_per_gpu = WORKLOADS[args.workload]
    w = make_workload(args.workload, dev, rank, torch, db)
    step, result, combine = w['step'], w['result'], w['combine']
    combine_desc = 'none'
    if dist is not None and combine is not None:
        combine_desc = 'NCCL allreduce of the per-GPU partials'
        if args.combine == 'peer':
            pc = sharding.enable_peer_combine()
            if pc.supports(result, sharding.resolve('Sum')):
                combine_desc = ('one-shot all-reduce kernel over NVLink peer memory (dace_b200_peer_allreduce), '
                                '%d B per GPU' % (result.numel() * result.element_size()))
    input_bytes = sum(t.numel() * t.element_size() for t in w['inputs'])
    needs_flush = input_bytes < (256 << 20)       # inputs smaller than ~2x L2: flush between iterations
    flush_buf = torch.empty(512 << 20, dtype=torch.uint8, device=dev) if needs_flush else None

    def full_step():
        step()
        if dist is not None and combine is not None:
            sharding.combine(result, 'Sum')   # peer-memory kernel or NCCL, same stream as the local kernel

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    use_graph = bool(w.get('graph')) and world == 1

    def timed(fn, k, graph=None):
        """Average device time per call (ms) over k calls, CUDA events on the launching stream."""
        if (use_graph if graph is None else graph) and not needs_flush:
            # launch-bound step: capture the k launches once, time one replay of the whole graph
            graph = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.graph(graph, stream=side):
                for _ in range(k):
                    fn()
            torch.cuda.synchronize()
            graph.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            graph.replay()
            e1.record()
            e1.synchronize()
            return e0.elapsed_time(e1) / k
        if not needs_flush:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(k):
                fn()
            e1.record()
            e1.synchronize()
            return e0.elapsed_time(e1) / k
        tot = 0.0
        for _ in range(k):
            flush_buf.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / k

    for _ in range(args.warmup):
        full_step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.05)
    barrier()
    ms = timed(full_step, args.steps)
    barrier()
    # kernel-only (no collective) timing for the roofline of the dominant kernel
    kernel_ms = timed(step, args.steps)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([ms, kernel_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, kernel_ms = t.tolist()
    value = world * bytes_per_gpu / (ms * 1e-3) / 1e9

    # ---- end to end: host buffers, H2D + kernel + D2H of the result inside the timed region ----
    e2e = None
    if args.e2e_steps > 0:
        e2e_inputs = w.get('e2e_inputs', w['inputs'])
        hosts = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in e2e_inputs]
        for h, t in zip(hosts, e2e_inputs):
            h.copy_(t)
        host_out = torch.empty(result.shape, dtype=result.dtype, pin_memory=True)

        e2e_prepare = w.get('e2e_prepare', lambda: None)

        def e2e_step():
            e2e_prepare()
            for h, t in zip(hosts, e2e_inputs):
                t.copy_(h, non_blocking=True)
            full_step()
            host_out.copy_(result, non_blocking=True)
            torch.cuda.current_stream().synchronize()

        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            e2e_step()
        barrier()
        e2e_s = (time.perf_counter() - t0) / args.e2e_steps
        if dist is not None:
            t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_s = t.item()
        e2e = {'value': world * bytes_per_gpu / e2e_s / 1e9, 'unit': UNIT, 'h2d_bytes_per_step': sum(t.numel() * t.element_size() for t in e2e_inputs),
               'd2h_bytes_per_step': result.numel() * result.element_size(), 'ms_per_step': e2e_s * 1e3,
               'steps': args.e2e_steps}
        del hosts

    launches_per_step = int(w.get('launches', 1)) + (1 if (dist is not None and combine is not None) else 0)
    l2_desc = w.get('l2') or ('L2 flushed (512 MiB write) between timed iterations' if needs_flush
                              else 'inputs (>= 1 GiB) far larger than the 126 MB L2')
    others = None
    if args.extras and world == 1:
        others = {}
        del w, step
        torch.cuda.empty_cache()
        for name in ('c1', 'c3', 'c3max', 'c4', 'c5'):
            if name == args.workload:
                continue
            ww = make_workload(name, dev, rank, torch, db)
            ib = sum(t.numel() * t.element_size() for t in ww['inputs'])
            needs_flush = ib < (256 << 20)
            if needs_flush and flush_buf is None:
                flush_buf = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
            for _ in range(3):
                ww['step']()
            torch.cuda.synchronize()
            t_ms = timed(ww['step'], args.steps, graph=bool(ww.get('graph')))
            others[name] = {'GB/s': WORKLOADS[name][1] / (t_ms * 1e-3) / 1e9, 'ms': t_ms,
                            'frac_of_hbm_peak': WORKLOADS[name][1] / (t_ms * 1e-3) / 1e9 / peaks()[0]}
            del ww
            torch.cuda.empty_cache()
        flush_buf = None
        torch.cuda.empty_cache()
        try:
            refgpu = reference_gpu_run(torch, dev)
        except Exception as ex:
            refgpu = {'error': str(ex)[:200]}
        if refgpu:
            others['reference_gpu'] = refgpu

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            cpu = cpu_reference_run(args.workload)
            cpu = {k: cpu[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')}
        except Exception as ex:   # baseline only; never blocks the GPU line
            cpu = {'value': None, 'unit': UNIT, 'cores': 0, 'kind': 'unavailable', 'sample': str(ex)}

    if rank == 0:
        peak, peak_src = peaks()
        achieved = bytes_per_gpu / (kernel_ms * 1e-3) / 1e9
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': dtype_of(args.workload), 'data': 'synthetic',
            'config': {'workload': desc, 'bytes_per_step_per_gpu': bytes_per_gpu,
                       'sharding': 'outer dimension, one shard per GPU' if world > 1 else 'single GPU',
                       'combine': combine_desc,
                       'l2': l2_desc},
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': UNIT, 'frac': achieved / peak,
                         'traffic': recorded_traffic(args.workload) if world == 1 else None, 'peak_source': peak_src, 'kernel_ms': kernel_ms,
                         'frac_of_8TBs_nominal': achieved / 8000.0},
            'cpu_baseline': cpu, 'e2e': e2e, 'gpu_launches': args.steps * launches_per_step, 'clocks': clocks,
        }
        if others is not None:
            line['others'] = others
        print(json.dumps(line))
    if dist is not None:
        sharding.disable_peer_combine()
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == '__main__':
    sys.exit(main())
