"""Per-rank launcher for sharded DaCe programs: ``python -m dace_b200.launch --nproc N script.py [args...]``.

The reference runs its distributed programs under ``mpirun`` (reference tests/conftest.py:66-96,
dace/sdfg/utils.py:1675-1708); this image has no MPI, and the generated programs of this backend need
nothing from it: one process per GPU, each seeing ITS GPU as device 0 (DaCe programs always select
device 0, reference dace/codegen/targets/cuda.py:446-455), a rank / world size, and a place to exchange
the 128-byte NCCL id (``dace_b200_comm_create_from_env``, include/dace_b200_reduce.h).  This launcher
provides exactly that and nothing else:

    RANK, WORLD_SIZE, LOCAL_RANK            as torchrun / mpirun wrappers set them
    CUDA_VISIBLE_DEVICES=<gpu of the rank>  (``--gpus 0,1,2,3`` to choose; default 0..N-1;
                                             ``--no-isolate`` leaves visibility alone)
    DACE_B200_RENDEZVOUS=<fresh temp file>  unique per launch, removed afterwards
    MASTER_ADDR / MASTER_PORT               127.0.0.1 and a free port, for scripts that also use torch.distributed

It waits for all ranks; if one fails the others are terminated (by PID) and the first non-zero exit
code is returned.
"""
import argparse
import os
import signal
import socket
import subprocess
import sys
import tempfile
import time
from typing import Dict, List, Optional, Sequence


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def rank_environment(rank: int, world: int, gpus: Optional[Sequence[int]], rendezvous: str, port: int,
                     base: Optional[Dict[str, str]] = None) -> Dict[str, str]:
    """Environment of one rank (pure function: the launcher's contract, unit-tested on the CPU)."""
    env = dict(os.environ if base is None else base)
    env.update({'RANK': str(rank), 'WORLD_SIZE': str(world), 'LOCAL_RANK': str(rank), 'LOCAL_WORLD_SIZE': str(world),
                'MASTER_ADDR': '127.0.0.1', 'MASTER_PORT': str(port), 'DACE_B200_RENDEZVOUS': rendezvous})
    if gpus is not None:
        env['CUDA_VISIBLE_DEVICES'] = str(gpus[rank])
    return env


def launch(argv: List[str], nproc: int, gpus: Optional[Sequence[int]] = None, isolate: bool = True,
           timeout: Optional[float] = None) -> int:
    if nproc < 1:
        raise ValueError('nproc must be >= 1')
    if isolate:
        gpus = list(gpus) if gpus is not None else list(range(nproc))
        if len(gpus) < nproc:
            raise ValueError('%d ranks but only %d GPUs listed' % (nproc, len(gpus)))
    else:
        gpus = None
    fd, rendezvous = tempfile.mkstemp(prefix='dace_b200_nccl_', suffix='.id')
    os.close(fd)
    os.unlink(rendezvous)          # rank 0 creates it (atomically) once it has an id
    port = _free_port()
    procs = []
    try:
        for r in range(nproc):
            procs.append(subprocess.Popen(argv, env=rank_environment(r, nproc, gpus, rendezvous, port)))
        deadline = None if timeout is None else time.monotonic() + timeout
        rc = 0
        pending = set(range(nproc))
        while pending:
            for r in sorted(pending):
                code = procs[r].poll()
                if code is None:
                    continue
                pending.discard(r)
                if code != 0 and rc == 0:
                    rc = code
            if rc != 0 or (deadline is not None and time.monotonic() > deadline):
                if rc == 0:
                    rc = 124
                break
            if pending:
                time.sleep(0.05)
        return rc
    finally:
        for p in procs:                      # only the exact PIDs started here
            if p.poll() is None:
                p.send_signal(signal.SIGTERM)
        for p in procs:
            try:
                p.wait(timeout=10)
            except subprocess.TimeoutExpired:
                p.kill()
                p.wait()
        for path in (rendezvous, rendezvous + '.tmp'):
            try:
                os.unlink(path)
            except FileNotFoundError:
                pass


def main(args: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(prog='python -m dace_b200.launch', description=__doc__.split('\n\n')[0])
    ap.add_argument('--nproc', type=int, required=True, help='number of ranks = number of GPUs used')
    ap.add_argument('--gpus', type=str, default=None, help='comma-separated GPU ordinals, one per rank (default 0..N-1)')
    ap.add_argument('--no-isolate', action='store_true', help='do not set CUDA_VISIBLE_DEVICES per rank')
    ap.add_argument('--timeout', type=float, default=None, help='seconds before all ranks are terminated')
    ap.add_argument('script', help='Python script (run with this interpreter) or executable')
    ap.add_argument('script_args', nargs=argparse.REMAINDER)
    ns = ap.parse_args(args)
    gpus = [int(g) for g in ns.gpus.split(',')] if ns.gpus else None
    cmd = ([sys.executable, ns.script] if ns.script.endswith('.py') else [ns.script]) + ns.script_args
    return launch(cmd, ns.nproc, gpus, not ns.no_isolate, ns.timeout)


if __name__ == '__main__':
    sys.exit(main())
