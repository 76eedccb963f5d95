#!/usr/bin/env python
"""Benchmark of the Reduce / WCR hot path on B200 (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c3max|c4|c5|c1] [--impl reference]

One "step" = one pass of the hot path over one batch of synthetic input already
resident in HBM.  Default workload at every N is BASELINE config C2 (full sum of
2^30 float32 to a scalar, per GPU); with N > 1 every rank reduces its own shard
(outer-dimension sharding) and the per-GPU partials are combined on the same
stream (one-shot NVLink peer-memory kernel, or NCCL) - weak scaling.  One JSON line
is printed by rank 0.  The default line also carries, under "others", the other
BASELINE configs (C1, C3, C3 single max, C4 with its key distributions, C5) timed on
the device, each with its roofline fraction and a parity check, and the reference's
own GPU expansions (CUB / GPUAuto / generated WCR kernel, built for sm_100 into
oracle/_ref) on the same shapes; at N > 1 "others" holds C5, the config that names
the 8-GPU split (16 KiB combine).  Every timed result is compared with an fp64
reduction of the same device inputs ("parity_ok").
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


def config_for(workload, world, combine='peer', dist_name='uniform'):
    """The `config` object of the JSON line - built the same way by both arms (--impl b200 / reference)."""
    desc, nbytes = WORKLOADS[workload]
    if workload == 'c4' and dist_name != 'uniform':
        desc = desc.replace('(uniform keys)', '(%s keys)' % dist_name)
    payload = {'c2': 4, 'c4': 4 << 20, 'c5': 16384}.get(workload)
    if world == 1 or payload is None:
        comb = 'none'
    elif combine == 'peer' and payload <= (64 << 10):
        comb = 'one-shot all-reduce kernel over NVLink peer memory (dace_b200_peer_allreduce), %d B per GPU' % payload
    else:
        comb = 'NCCL allreduce of the per-GPU partials, %d B per GPU' % payload
    if workload == 'c1':
        l2 = ('ring of 16 distinct 4096x4096 inputs (1 GiB, 8x the 126 MB L2): consecutive launches read '
              'different matrices, no flush needed; the K timed launches are replayed from one CUDA graph '
              '(a ~14 us kernel is otherwise bound by the Python call overhead)')
    else:
        l2 = 'inputs (>= 1 GiB) far larger than the 126 MB L2'
    return {'workload': desc, 'bytes_per_step_per_gpu': nbytes,
            'sharding': 'outer dimension, one shard per GPU' if world > 1 else 'single GPU',
            'combine': comb, 'l2': l2}


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


def make_workload(name, dev, rank, torch, db, dist_name='uniform'):
    """Inputs on the device + `step` (one pass of the hot path) + `check` (fp64 reduction of the same
    device inputs -> (expected tensor, tolerance rule)) for one BASELINE config."""
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
        w['expect'] = lambda: torch.sum(x, dtype=torch.float64).reshape(())
        w['rule'] = 'rel1e-5'
    elif name in ('c3', 'c3max'):
        x = torch.randn(65536, 8192, dtype=torch.float32, device=dev, generator=g).to(torch.bfloat16)
        mx = torch.empty(65536, dtype=torch.float32, device=dev)
        sm = torch.empty(65536, dtype=torch.float32, device=dev)
        w['inputs'] = [x]
        if name == 'c3':
            w['step'] = lambda: db.rows_max_sum(x, mx, sm)
            w['extra_check'] = lambda: bool(
                torch.all((sm.double() - torch.sum(x, dim=1, dtype=torch.float64)).abs()
                          <= 1e-5 * torch.sum(x.abs(), dim=1, dtype=torch.float64)).item())
        else:
            w['step'] = lambda: db.reduce(MAX, x, mx, axis=1, identity=-3.0e38)
        w['result'] = mx
        w['combine'] = None   # rows are sharded along the kept axis: no collective
        w['expect'] = lambda: torch.amax(x, dim=1).double()
        w['rule'] = 'exact'
    elif name == 'c4':
        n, nb = 1 << 28, 1 << 20
        if dist_name == 'uniform':
            keys = torch.randint(0, nb, (n, ), dtype=torch.int32, device=dev, generator=g)
        elif dist_name == 'all_equal':
            keys = torch.full((n, ), 12345, dtype=torch.int32, device=dev)
        elif dist_name == 'zipf1.1':
            # discrete power law with exponent 1.1 by inversion of the continuous CDF (P(K >= k) = k^-0.1),
            # wrapped into the bin range like tools/c4_probe.py: ~13 % of the keys hit one bin
            u = torch.rand(n, dtype=torch.float64, device=dev, generator=g).clamp_(min=1e-12)
            k = torch.pow(u, -10.0).clamp_(max=2.0 ** 62).floor_().to(torch.int64)
            keys = (k % nb).to(torch.int32)
            del u, k
        else:
            raise SystemExit('unknown key distribution %r' % dist_name)
        bins = torch.zeros(nb, dtype=torch.int32, device=dev)
        w['inputs'] = [keys]

        def step():
            bins.zero_()          # part of the step: WCR accumulates into its output
            db.wcr_scatter(ADD, bins, keys)
        w['step'] = step
        w['result'] = bins
        w['combine'] = 'sum'
        w['launches'] = 3   # key sampler + partitioned kernel + gated global-atomic kernel (torch's zero_ not counted)
        w['expect'] = lambda: torch.bincount(keys, minlength=nb).double()
        w['rule'] = 'exact'
    elif name == 'c5':
        x = torch.rand(1 << 19, 4096, dtype=torch.float32, device=dev, generator=g)
        out = torch.zeros(4096, dtype=torch.float32, device=dev)
        w['inputs'] = [x]
        w['step'] = lambda: db.reduce(ADD, x, out, axis=0, identity=0)
        w['result'] = out
        w['combine'] = 'sum'
        w['expect'] = lambda: torch.sum(x, dim=0, dtype=torch.float64)
        w['rule'] = 'rel1e-5'
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
        w['graph'] = True
        # after any number of steps `out` holds the row sums of the matrix read last
        w['expect'] = lambda: torch.sum(ring[(pos[0] - 1) & 15], dim=1, dtype=torch.float64)
        w['rule'] = 'rel1e-5'
    else:
        raise SystemExit('unknown workload %r' % name)
    return w


def parity_ok(w, torch, dist=None, combined=False):
    """Compare the result the timed loop left behind with an fp64 reduction of the same device inputs
    (N > 1 with a combine: with the all-reduced expectation).  Sums: <= 1e-5 relative to the result
    (all inputs are non-negative); max / histogram counts: exact."""
    got = w['result'].detach().double()
    exp = w['expect']()
    if combined and dist is not None and w['combine'] is not None:
        dist.all_reduce(exp, op=dist.ReduceOp.SUM)
    if w['rule'] == 'exact':
        ok = bool(torch.equal(got.reshape(exp.shape), exp))
    else:
        ok = bool(torch.all((got.reshape(exp.shape) - exp).abs() <= 1e-5 * exp.abs()).item())
    if 'extra_check' in w and not combined:
        ok = ok and bool(w['extra_check']())
    if dist is not None:
        t = torch.tensor([1 if ok else 0], device=got.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        ok = bool(t.item())
    return ok


# ---------------------------------------------------------------------------
# CPU side: the reference's own CPU expansion (oracle/_ref, built from the
# reference by oracle/build_ref.py); the C oracle port if the reference could
# not be compiled.  Checker/baseline only.
# ---------------------------------------------------------------------------
def _set_omp_threads(cores):
    """torchrun exports OMP_NUM_THREADS=1 to its workers and libgomp may already be initialised in this
    process: set the thread count through the runtime itself and report what it says."""
    import ctypes
    os.environ['OMP_NUM_THREADS'] = str(cores)
    try:
        gomp = ctypes.CDLL('libgomp.so.1', mode=ctypes.RTLD_GLOBAL)
        gomp.omp_set_num_threads(int(cores))
        gomp.omp_get_max_threads.restype = ctypes.c_int
        return int(gomp.omp_get_max_threads())
    except OSError:
        return None


def cpu_reference_prepare(workload):
    """Host inputs of the FULL per-GPU workload (same shapes as the GPU arm) + the reference program."""
    import numpy as np
    from oracle import ref_loader
    cores = os.cpu_count() or 1
    threads = _set_omp_threads(cores)
    rng = np.random.default_rng(0)
    if workload == 'c2':
        n = 1 << 30
        sample = 'all 2^30 float32 (4 GiB, the full config), reference OpenMP expansion'
        data = {'A': rng.random(n, dtype=np.float32), 'B': np.zeros(1, np.float32), 'D0': n}
        prog, nbytes = 'full_sum_f32_omp', WORKLOADS['c2'][1]
    elif workload in ('c3', 'c3max'):
        rows, cols = 8192, 8192
        sample = '[8192, 8192] bf16 slice (128 MiB of the 1 GiB), reference pure expansion (max, then sum for c3)'
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
        sample = '[2^16, 4096] float32 slice (1 GiB of the 8 GiB shard: ~1 s per pass), reference OpenMP expansion'
        data = {'A': rng.random((rows, cols), dtype=np.float32), 'B': np.zeros(cols, np.float32), 'D0': rows, 'D1': cols}
        prog, nbytes = 'col_sum_f32_omp', rows * cols * 4
    else:
        rows, cols = 4096, 4096
        sample = 'full 4096x4096 float32, reference @dace.program np.sum(A, axis=1), OpenMP expansion'
        data = {'A': rng.random((rows, cols), dtype=np.float32), 'B': np.zeros(rows, np.float32), 'N': rows, 'M': cols}
        prog, nbytes = 'c1_npsum_axis1_f32_omp', WORKLOADS['c1'][1]

    if ref_loader.available():
        kind = 'reference'
        p = ref_loader.RefProgram(prog)
        extra = ref_loader.RefProgram('row_sum_bf16_f32_pure') if workload == 'c3' else None

        def run():
            p(**data)
            if extra is not None:
                extra(**data)
    else:
        kind, cores, threads = 'port', 1, 1
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

    def check():
        """The baseline's own result against NumPy fp64: (ok at 1e-5 relative, worst relative error).  The
        reference's fp32 OpenMP full sum accumulates each thread's slice sequentially in fp32 and stalls at
        2^24 per thread on C2's 2^30 uniform [0,1) values - its error is reported, not hidden."""
        if workload == 'c2':
            ref = np.array([np.sum(data['A'], dtype=np.float64)])
        elif workload in ('c5', 'c1'):
            ref = np.sum(data['A'], axis=0 if workload == 'c5' else 1, dtype=np.float64)
        else:
            return None, None
        err = float(np.max(np.abs(data['B'].astype(np.float64) - ref) / np.abs(ref)))
        return err <= 1e-5, err

    return {'run': run, 'check': check, 'nbytes': nbytes, 'sample': sample, 'kind': kind,
            'cores': threads or cores, 'omp_threads': threads}


def cpu_reference_run(workload, budget_s=12.0, max_reps=5, prepared=None):
    c = prepared or cpu_reference_prepare(workload)
    c['run']()  # warm-up (page faults, OpenMP pool)
    times = []
    t_start = time.perf_counter()
    while len(times) < max_reps and (time.perf_counter() - t_start < budget_s or not times):
        t0 = time.perf_counter()
        c['run']()
        times.append(time.perf_counter() - t0)
    best = min(times)
    return {'value': c['nbytes'] / best / 1e9, 'unit': UNIT, 'cores': c['cores'], 'kind': c['kind'],
            'sample': c['sample'], 'best_s': best, 'reps': len(times), 'nbytes': c['nbytes'],
            'omp_threads': c['omp_threads'], 'parity_ok': c['check']()[0], 'rel_err_vs_fp64': c['check']()[1]}


# ---------------------------------------------------------------------------
# The reference's own GPU expansions (CUB 'CUDA (device)', 'GPUAuto', generated
# WCR-map kernel) built for sm_100 by oracle/build_ref.py, on device-resident
# arrays of the same shapes: the on-box GPU comparison.  The generated programs
# launch on their own streams, so they are timed by wall clock with a device
# synchronise after every call.
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
    """--impl reference: the reference's own CPU expansion of the path on the host cores, all threads,
    on the full per-GPU workload where it fits a step of a fraction of a second (C2, C1), else a stated slice."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    prepared = cpu_reference_prepare(args.workload)
    res = []
    for _ in range(max(1, args.warmup)):
        cpu_reference_run(args.workload, budget_s=0.0, max_reps=1, prepared=prepared)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res.append(cpu_reference_run(args.workload, budget_s=0.0, max_reps=1, prepared=prepared))
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
        'config': config_for(args.workload, args.gpus, args.combine, args.dist),
        'cpu_baseline': {k: base[k] for k in ('value', 'unit', 'cores', 'kind', 'sample', 'omp_threads')},
        'e2e': {'value': best, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'parity_ok': base['parity_ok'], 'rel_err_vs_fp64': base['rel_err_vs_fp64'],
    }
    print(json.dumps(line))
    return 0


def dtype_of(workload):
    return {'c2': 'f32', 'c3': 'bf16->f32', 'c3max': 'bf16->f32', 'c4': 'i32', 'c5': 'f32', 'c1': 'f32'}[workload]


BUILT_DIR = os.path.join(ROOT, 'tests', 'dace_e2e', '_built')


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
    ap.add_argument('--extras', action='store_true', help='(default now) time the other configs into "others"')
    ap.add_argument('--no-extras', action='store_true', help='skip the "others" table')
    ap.add_argument('--dist', default='uniform', choices=('uniform', 'zipf1.1', 'all_equal'),
                    help='key distribution of the c4 workload')
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
    w = make_workload(args.workload, dev, rank, torch, db, args.dist)
    step, result, combine = w['step'], w['result'], w['combine']
    peer = None
    if dist is not None and args.combine == 'peer':
        peer = sharding.enable_peer_combine()
    flush = {'buf': None}

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def make_full_step(ww):
        def full_step():
            ww['step']()
            if dist is not None and ww['combine'] is not None:
                sharding.combine(ww['result'], 'Sum')   # peer-memory kernel or NCCL, same stream as the local kernel
        return full_step

    def timed(fn, k, needs_flush, graph):
        """Average device time per call (ms) over k calls, CUDA events on the launching stream."""
        if graph and not needs_flush:
            # launch-bound step: capture the k launches once, time one replay of the whole graph
            cg_ = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.graph(cg_, stream=side):
                for _ in range(k):
                    fn()
            torch.cuda.synchronize()
            cg_.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            cg_.replay()
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
        if flush['buf'] is None:
            flush['buf'] = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
        tot = 0.0
        for _ in range(k):
            flush['buf'].fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / k

    def measure(ww, k, warmup):
        """(ms per full step, ms per kernel-only step, parity) for one workload, max over ranks."""
        ib = sum(t.numel() * t.element_size() for t in ww['inputs'])
        needs_flush = ib < (256 << 20)       # inputs smaller than ~2x L2: flush between iterations
        graph = bool(ww.get('graph')) and world == 1
        fs = make_full_step(ww)
        for _ in range(warmup):
            fs()
        barrier()
        ms_full = timed(fs, k, needs_flush, graph)
        barrier()
        ok_combined = parity_ok(ww, torch, dist, combined=True)
        ms_kernel = timed(ww['step'], k, needs_flush, graph)
        barrier()
        ok_local = parity_ok(ww, torch, dist, combined=False)
        if dist is not None:
            t = torch.tensor([ms_full, ms_kernel], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_full, ms_kernel = t.tolist()
        return ms_full, ms_kernel, ok_combined and ok_local

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.05)
    ms, kernel_ms, ok = measure(w, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    value = world * bytes_per_gpu / (ms * 1e-3) / 1e9
    full_step = make_full_step(w)

    # ---- end to end: host buffers, H2D + kernel + D2H of the result inside the timed region ----
    e2e = None
    e2e_resident = None
    if args.e2e_steps > 0:
        e2e_inputs = w.get('e2e_inputs', w['inputs'])
        hosts = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in e2e_inputs]
        for h, t in zip(hosts, e2e_inputs):
            h.copy_(t)
        host_out = torch.empty(result.shape, dtype=result.dtype, pin_memory=True)
        h2d_bytes = sum(t.numel() * t.element_size() for t in e2e_inputs)
        e2e_prepare = w.get('e2e_prepare', lambda: None)

        # the link's own speed: the same pinned buffers copied with nothing else in the step (all ranks at once:
        # the host side of the link is shared)
        barrier()
        t0 = time.perf_counter()
        for h, t in zip(hosts, e2e_inputs):
            t.copy_(h, non_blocking=True)
        torch.cuda.synchronize()
        h2d_gbs = h2d_bytes / (time.perf_counter() - t0) / 1e9

        from dace_b200 import compiled
        plugin = None
        if world == 1 and args.workload == 'c2':
            for cand in ('b200_full_sum_f32_persistent', 'b200_full_sum_f32'):
                if compiled.available(BUILT_DIR, cand):
                    plugin = compiled.CompiledProgram(BUILT_DIR, cand)
                    break
        if plugin is not None:
            # Through the plug-in: __program_<name> of the DaCe-built .so (Reduce node expanded with 'B200'),
            # host arrays in and out - the generated code does H2D, calls dace_b200_reduce_nd and copies back,
            # the way CompiledSDFG.fast_call runs it (reference dace/codegen/compiled_sdfg.py:549-592).
            call = plugin.bind(A=hosts[0], B=host_out.reshape(1), D0=hosts[0].numel())
            path = '__program_%s of %s (DaCe-generated, ctypes)' % (plugin.name, os.path.relpath(plugin.path, ROOT))

            def e2e_step():
                call()
                torch.cuda.synchronize()
        else:
            path = 'dace_b200.reduce host API on torch tensors (H2D copy_, kernel, combine, D2H copy_)'

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
        e2e_ok = None
        if plugin is not None:
            exp = float(w['expect']().item())
            e2e_ok = abs(float(host_out.reshape(-1)[0]) - exp) <= 1e-5 * abs(exp)
        if dist is not None:
            t = torch.tensor([e2e_s, -h2d_gbs], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_s, h2d_gbs = t[0].item(), -t[1].item()
        e2e_val = world * bytes_per_gpu / e2e_s / 1e9
        e2e = {'value': e2e_val, 'unit': UNIT, 'h2d_bytes_per_step': h2d_bytes,
               'd2h_bytes_per_step': result.numel() * result.element_size(), 'ms_per_step': e2e_s * 1e3,
               'steps': args.e2e_steps, 'path': path, 'parity_ok': e2e_ok,
               'roofline': {'bound': 'pcie_h2d', 'achieved': e2e_val / world, 'peak': h2d_gbs, 'unit': UNIT,
                            'frac': e2e_val / world / h2d_gbs,
                            'peak_source': 'cudaMemcpyAsync of the same pinned buffers alone, measured in this run '
                                           '(per GPU, slowest rank)'},
               'numa': 'pinned host buffers are allocated by the rank that uses them; the 8-GPU boxes of this pool have '
                       'ONE NUMA node (profiles/r2_numa_probe_8gpu.txt: node0 = cpus 0-31, every GPU NUMA affinity 0), so '
                       'there is no placement to choose; 8 ranks copying at once share 231 GB/s of host-side bandwidth '
                       '(4 GPUs at 23 GB/s, 4 at 35 GB/s)'}
        if plugin is not None:
            plugin.close()
        # device-resident arrays through the plug-in: launch + the program's own stream synchronise
        if world == 1 and args.workload == 'c2' and compiled.available(BUILT_DIR, 'b200_full_sum_f32_resident'):
            rp = compiled.CompiledProgram(BUILT_DIR, 'b200_full_sum_f32_resident')
            rout = torch.zeros(1, dtype=torch.float32, device=dev)
            rcall = rp.bind(A=w['inputs'][0], B=rout, D0=w['inputs'][0].numel())
            rcall()
            torch.cuda.synchronize()
            k = max(10, min(args.steps, 50))
            t0 = time.perf_counter()
            for _ in range(k):
                rcall()
            torch.cuda.synchronize()
            r_s = (time.perf_counter() - t0) / k
            exp = float(w['expect']().item())
            e2e_resident = {'value': bytes_per_gpu / r_s / 1e9, 'unit': UNIT, 'ms_per_step': r_s * 1e3, 'steps': k,
                            'path': '__program_b200_full_sum_f32_resident (GPU_Global arrays, wall clock incl. ctypes '
                                    'call and the generated cudaStreamSynchronize)',
                            'parity_ok': abs(float(rout.item()) - exp) <= 1e-5 * abs(exp)}
            rp.close()
        del hosts

    launches_per_step = int(w.get('launches', 1)) + (1 if (dist is not None and combine is not None) else 0)
    peak, peak_src = peaks()

    def entry(name, ww, k, traffic_key=None):
        m_full, m_kernel, okk = measure(ww, k, 3)
        nb = WORKLOADS[name][1]
        ach = nb / (m_kernel * 1e-3) / 1e9
        e = {'GB/s': world * nb / (m_full * 1e-3) / 1e9, 'ms_per_step': m_full, 'kernel_ms': m_kernel,
             'frac_of_hbm_peak': ach / peak, 'traffic': recorded_traffic(traffic_key or name) if world == 1 else None,
             'parity_ok': okk}
        if world > 1 and ww['combine'] is not None:
            e['combine_us'] = (m_full - m_kernel) * 1e3
        return e

    others = None
    if not args.no_extras:
        others = {}
        del w, step, full_step
        torch.cuda.empty_cache()
        names = ('c1', 'c3', 'c3max', 'c4', 'c5') if world == 1 else ('c5', )
        for name in names:
            if name == args.workload:
                continue
            ww = make_workload(name, dev, rank, torch, db)
            others[name] = entry(name, ww, min(args.steps, 50))
            del ww
            torch.cuda.empty_cache()
        if world == 1 and args.workload != 'c4':
            for dn in ('zipf1.1', 'all_equal'):
                ww = make_workload('c4', dev, rank, torch, db, dn)
                others['c4_' + dn.replace('.', '_')] = entry('c4', ww, 10, 'c4_' + dn.replace('.', '_'))
                del ww
                torch.cuda.empty_cache()
        flush['buf'] = None
        torch.cuda.empty_cache()
        if world == 1:
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
            cpu = {k: cpu[k] for k in ('value', 'unit', 'cores', 'kind', 'sample', 'omp_threads', 'parity_ok',
                                       'rel_err_vs_fp64')}
        except Exception as ex:   # baseline only; never blocks the GPU line
            cpu = {'value': None, 'unit': UNIT, 'cores': 0, 'kind': 'unavailable', 'sample': str(ex)}

    if rank == 0:
        achieved = bytes_per_gpu / (kernel_ms * 1e-3) / 1e9
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': dtype_of(args.workload), 'data': 'synthetic',
            'config': config_for(args.workload, world, args.combine if peer is not None or world == 1 else 'nccl', args.dist),
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': UNIT, 'frac': achieved / peak,
                         'traffic': recorded_traffic(args.workload if args.dist == 'uniform' else 'c4_' + args.dist.replace('.', '_')) if world == 1 else None,
                         'peak_source': peak_src, 'kernel_ms': kernel_ms,
                         'frac_of_8TBs_nominal': achieved / 8000.0},
            'parity_ok': ok,
            'cpu_baseline': cpu, 'e2e': e2e, 'gpu_launches': args.steps * launches_per_step, 'clocks': clocks,
        }
        if world > 1 and combine is not None:
            line['combine_us'] = (ms - kernel_ms) * 1e3
        if e2e_resident is not None:
            line['e2e_resident'] = e2e_resident
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
