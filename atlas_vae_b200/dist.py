"""One process per GPU (torchrun); NCCL over NVLink for the gradient all-reduce, gloo on CPU-only test boxes.

The reference's only data-parallel semantics are jet-ID's tf.distribute.MirroredStrategy
(jet-ID/models.py:69-73, classifier.py:136-138): synchronous replicas, gradients summed.  Here every rank
owns a contiguous slice of each global batch and the flat fp32 gradient is all-reduced with SUM (the reference
differentiates the SUM of per-jet losses, models.py:183), so n ranks reproduce the 1-rank update.
"""
import os

import torch
import torch.distributed as td


def rank_world():
    if td.is_available() and td.is_initialized():
        return td.get_rank(), td.get_world_size()
    return 0, 1


def init(backend=None):
    """Initialise torch.distributed from the torchrun environment (no-op for single-process runs)."""
    world = int(os.environ.get('WORLD_SIZE', '1'))
    if world == 1 or td.is_initialized():
        return rank_world()
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    os.environ.setdefault('MASTER_PORT', '29500')
    if backend is None:
        backend = 'nccl' if torch.cuda.is_available() else 'gloo'
    if backend == 'nccl':
        torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', '0')))
    td.init_process_group(backend=backend, rank=int(os.environ['RANK']), world_size=world)
    return rank_world()


def all_reduce_sum(t):
    if td.is_available() and td.is_initialized() and td.get_world_size() > 1:
        td.all_reduce(t, op=td.ReduceOp.SUM)
    return t


def all_reduce_minmax(lo, hi):
    if td.is_available() and td.is_initialized() and td.get_world_size() > 1:
        td.all_reduce(lo, op=td.ReduceOp.MIN)
        td.all_reduce(hi, op=td.ReduceOp.MAX)
    return lo, hi


def shard_bounds(n, rank, world):
    """Contiguous shard [lo, hi) of n units for `rank` (scoring: SURVEY 8(e))."""
    return (n * rank) // world, (n * (rank + 1)) // world


def barrier():
    if td.is_available() and td.is_initialized() and td.get_world_size() > 1:
        td.barrier()


def bind_to_gpu_numa(local_rank):
    """Pin this process to the CPUs of the NUMA node its GPU hangs off (sysfs), BEFORE it allocates host buffers, so that the
    page-locked batches a rank uploads are first-touched on the node next to its GPU.  Returns a short description (for bench
    records); does nothing when the topology is not exposed."""
    try:
        out = __import__('subprocess').run(['nvidia-smi', '--query-gpu=pci.bus_id', '--format=csv,noheader', '-i', str(local_rank)],
                                           capture_output=True, text=True, timeout=20).stdout.strip().lower()
        bus = out[-12:] if len(out) >= 12 else out                 # 00000000:1b:00.0 -> 0000:1b:00.0
        base = '/sys/bus/pci/devices/' + bus
        node = int(open(base + '/numa_node').read().strip())
        cl = open(base + '/local_cpulist').read().strip()
        cpus = set()
        for part in cl.split(','):
            a, _, b = part.partition('-')
            cpus |= set(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if node < 0 or not cpus:
            return 'gpu %s: no NUMA affinity exposed' % bus
        os.sched_setaffinity(0, cpus)
        return 'gpu %s -> NUMA node %d (%d cpus)' % (bus, node, len(cpus))
    except Exception as e:                                          # noqa: BLE001 -- diagnostics only
        return 'not bound (%s)' % type(e).__name__
