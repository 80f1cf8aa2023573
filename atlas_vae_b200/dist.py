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
