"""Data parallelism for the path: one process per GPU, gradients averaged across ranks.

Mirrors what the reference gets from `torch.nn.parallel.DistributedDataParallel`
(imaginaire/trainers/utils/get_trainer.py:81-88) and `init_dist` (imaginaire/utils/distributed.py:21-27),
with two changes that matter at this model's shape: (1) DDP all-reduces the whole 1.46 GB table gradient
every step even when only 4 of 16 levels are active -- here only the ACTIVE-LEVEL PREFIX of the table
gradient is reduced (levels >= active have identically zero gradient, modules.py:91-93,110-113); (2) all
other parameters live in one flat gradient bucket, so the step issues exactly two collectives.  The SUM is
taken by the collective and the 1/world_size factor is folded into the fused optimizer step.
"""
import os

import torch
import torch.distributed as dist


def init_dist(backend=None):
    """env:// rendezvous from torchrun (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*).  Returns (rank, local_rank, world)."""
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kwargs = {}
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            kwargs["device_id"] = torch.device("cuda", local_rank)
        dist.init_process_group(backend, **kwargs)
    return rank, local_rank, world


def world_size():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def broadcast_parameters(tensors, src=0):
    """Identical replicas at start (DDP broadcasts rank 0's parameters on construction)."""
    if world_size() > 1:
        for t in tensors:
            dist.broadcast(t, src=src)


def allreduce_gradients(flat_grad, table_grad_prefix):
    """SUM-all-reduce the two gradient buffers in place; returns the factor (1/world) still to be applied.
    The small bucket goes first so that it completes under the large transfer."""
    w = world_size()
    if w == 1:
        return 1.0
    handles = [dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, async_op=True)]
    if table_grad_prefix is not None and table_grad_prefix.numel() > 0:
        handles.append(dist.all_reduce(table_grad_prefix, op=dist.ReduceOp.SUM, async_op=True))
    for h in handles:
        h.wait()
    return 1.0 / w


def shard_indices(n_items, rank, world):
    """Round-robin sharding `idx = rank, rank + world, ...` (DistributedSampler(shuffle=False), mesh.py:104-105),
    without the sampler's wrap-around padding."""
    return list(range(rank, n_items, world))
