"""Synchronous data parallelism across the GPUs of one box: the replacement for the reference's
parameter server (reference src/main.py:100-193) with the semantics of its --sync mode
(SyncReplicasOptimizer: mean of the N replica gradients, one apply; src/network.py:187-192).

One process per GPU (torchrun).  Every rank owns its own replay ring -- the reference already gives
each worker its own ReplayMemory (main.py:151) -- samples a local minibatch, and computes the mean
gradient of the DATA term over that minibatch.  The only exchange is one all-reduce (sum) of the flat
gradient buffer; the update then uses grad_scale = 1/world (mean over ranks) and adds the L2 term once,
so an N-rank step on batches B_r equals a 1-rank step on concat(B_r).
"""
from __future__ import annotations

import os


def world_info():
    """(rank, local_rank, world_size) from torchrun's environment (1 process = 1 GPU)."""
    return (int(os.environ.get('RANK', '0')), int(os.environ.get('LOCAL_RANK', '0')),
            int(os.environ.get('WORLD_SIZE', '1')))


def init(backend=None, device=None):
    """Initialises torch.distributed from the environment if WORLD_SIZE > 1; returns world size."""
    import torch
    import torch.distributed as dist
    rank, local, world = world_info()
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = 'nccl' if torch.cuda.is_available() else 'gloo'
        kw = {}
        if backend == 'nccl' and device is not None:
            kw['device_id'] = device
        dist.init_process_group(backend, **kw)
    return world


def grad_scale(world):
    """Factor applied to the all-reduced (summed) gradient: the SyncReplicas mean."""
    return 1.0 / float(world)


def sync_gradients(flat_grad):
    """In-place sum of the flat gradient buffer over all ranks (NCCL over NVLink on GPUs)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(flat_grad)
    return flat_grad


def broadcast_parameters(flat_params, src=0):
    """Rank `src` initialises, everyone else copies (the reference's chief/non-chief start-up, main.py:133-143)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(flat_params, src)
    return flat_params


def shard_seed(base_seed, rank):
    """Per-rank seed so replay shards and sampling streams differ between ranks."""
    return None if base_seed is None else int(base_seed) + 1000003 * int(rank)
