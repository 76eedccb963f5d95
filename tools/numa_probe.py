"""H2D bandwidth of pinned host buffers under different NUMA memory policies, all ranks copying at once
(torchrun --nproc-per-node N tools/numa_probe.py).  Decides where bench.py's end-to-end leg allocates its
pinned buffers at N > 1 (e2e aggregate was 0.42 of linear at 8 GPUs with first-touch placement)."""
import ctypes, glob, os, sys, time
import torch
import torch.distributed as dist

MPOL_DEFAULT, MPOL_PREFERRED, MPOL_BIND, MPOL_INTERLEAVE = 0, 1, 2, 3
libc = ctypes.CDLL('libc.so.6', use_errno=True)


def set_mempolicy(mode, nodes):
    mask = 0
    for n in nodes:
        mask |= 1 << n
    m = ctypes.c_ulong(mask)
    rc = libc.syscall(238, mode, ctypes.byref(m), 64)      # __NR_set_mempolicy on x86-64
    if rc != 0:
        return 'errno %d' % ctypes.get_errno()
    return 'ok'


rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
if world > 1:
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
nodes = sorted(int(p.rsplit('node', 1)[1]) for p in glob.glob('/sys/devices/system/node/node[0-9]*'))
if rank == 0:
    print('numa nodes:', nodes, 'cpus:', os.cpu_count(), flush=True)
    for n in nodes:
        try:
            print(' node%d cpulist' % n, open('/sys/devices/system/node/node%d/cpulist' % n).read().strip(), flush=True)
        except OSError:
            pass
    os.system('nvidia-smi topo -m 2>/dev/null | head -14')
nbytes = 2 << 30
dst = torch.empty(nbytes, dtype=torch.uint8, device=dev)
policies = [('default (first touch)', MPOL_DEFAULT, [])] + [('bind node%d' % n, MPOL_BIND, [n]) for n in nodes]
if len(nodes) > 1:
    policies += [('interleave all', MPOL_INTERLEAVE, nodes), ('bind node(rank %% nodes)', MPOL_BIND, [nodes[rank % len(nodes)]])]
for name, mode, ns in policies:
    st = set_mempolicy(mode, ns)
    host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    host.fill_(1)
    set_mempolicy(MPOL_DEFAULT, [])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        dst.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 3
    gbs = torch.tensor([nbytes / dt / 1e9], dtype=torch.float64, device=dev)
    if world > 1:
        allg = [torch.zeros_like(gbs) for _ in range(world)]
        dist.all_gather(allg, gbs)
        vals = [a.item() for a in allg]
    else:
        vals = [gbs.item()]
    if rank == 0:
        print('%-28s (%s): per rank %s  aggregate %.1f GB/s' % (name, st, ' '.join('%.1f' % v for v in vals), sum(vals)), flush=True)
    del host
if world > 1:
    dist.destroy_process_group()
