"""Synchronous data parallelism across the GPUs of one box: the replacement for the reference's
parameter server (reference src/main.py:100-193) with the semantics of its --sync mode
(SyncReplicasOptimizer: mean of the N replica gradients, one apply; src/network.py:187-192).

One process per GPU (torchrun).  Every rank owns its own replay ring -- the reference already gives
each worker its own ReplayMemory (main.py:151) -- samples a local minibatch, and computes the mean
gradient of the DATA term over that minibatch.  The update uses grad_scale = 1/world (mean over
ranks) and adds the L2 term once, so an N-rank step on batches B_r equals a 1-rank step on concat(B_r).

Two exchanges (``setup_network(..., mode=)``):
* ``peer`` -- the gradient exchange is fused into the step's own kernels over NVLink peer memory
  (``Network.connect_peers`` / ``dp_train_step``; csrc/net.cu);
* ``nccl`` -- ``backward`` -> ``all_reduce`` of the flat gradient buffer -> ``apply``.

``main.py``, ``bench.py`` and the tests all come through this module.  ``StopAgreement`` keeps the number
of exchanges equal on every rank when the ranks' host loops (episodes of different length) finish
at different times.
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


def is_distributed():
    try:
        import torch.distributed as dist
        return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    except Exception:
        return False


def grad_scale(world):
    """Factor applied to the all-reduced (summed) gradient: the SyncReplicas mean."""
    return 1.0 / float(world)


def sync_gradients(flat_grad):
    """In-place sum of the flat gradient buffer over all ranks (NCCL over NVLink on GPUs)."""
    import torch.distributed as dist
    if is_distributed():
        dist.all_reduce(flat_grad)
    return flat_grad


def broadcast_parameters(flat_params, src=0):
    """Rank `src` initialises, everyone else copies (the reference's chief/non-chief start-up, main.py:133-143).
    For a Network use ``setup_network`` / ``Network.broadcast_parameters``, which also rebuild the bf16 operand copies."""
    import torch.distributed as dist
    if is_distributed():
        dist.broadcast(flat_params, src)
    return flat_params


def shard_seed(base_seed, rank):
    """Per-rank seed so replay shards and sampling streams differ between ranks."""
    return None if base_seed is None else int(base_seed) + 1000003 * int(rank)


def setup_network(network, config, mode='peer'):
    """Start-up of one replica: copy rank 0's parameters, mean-of-gradients scaling, and (mode 'peer') map the
    ranks' exchange blocks.  Returns the world size."""
    import torch.distributed as dist
    if not is_distributed():
        return 1
    world = dist.get_world_size()
    network.broadcast_parameters(0)
    network.set_hyper(config, grad_scale=grad_scale(world))
    if mode == 'peer':
        network.connect_peers()
    elif mode != 'nccl':
        raise RuntimeError('Unsupported data-parallel exchange {}'.format(mode))
    return world


def train_step_table(network, replay_memory, index_table, want_loss=False, loss_out=None):
    """``train_step`` on the current row of a device-resident ``replay_memory.IndexTable`` (the update advances the
    table's cursor): single GPU or peer-memory exchange; the NCCL path takes the row with ``IndexTable.next``."""
    if not is_distributed() or getattr(network, 'peers_connected', False):
        return network.train_step_table(replay_memory, index_table, want_loss=want_loss, loss_out=loss_out,
                                        data_parallel=is_distributed())
    return train_step(network, replay_memory, index_table.k, index_table.next(index_table.scratch), want_loss=want_loss,
                      loss_out=loss_out)


def train_step(network, replay_memory, k, d_indices, want_loss=False, loss_out=None):
    """One synchronous update from this rank's replay shard: single GPU, peer-memory exchange, or NCCL all-reduce,
    whichever the network was set up for.  Every rank must call it the same number of times."""
    if not is_distributed():
        return network.train_step(replay_memory, k, d_indices=d_indices, want_loss=want_loss, loss_out=loss_out)
    if getattr(network, 'peers_connected', False):
        return network.dp_train_step(replay_memory, k, d_indices=d_indices, want_loss=want_loss, loss_out=loss_out)
    batch = replay_memory.gather(d_indices)
    loss, _ = network.backward_on_batch(batch, want_loss=want_loss)
    sync_gradients(network.flat_buffer('grad'))                   # sum; grad_scale = 1/world applied by the update
    network.apply_gradients()
    if want_loss and loss_out is not None:
        loss_out.copy_(loss, non_blocking=True)
        return loss_out
    return loss


class StopAgreement(object):
    """Keeps the number of gradient exchanges rank-independent.

    The reference's workers stop whenever their own episode loop ends (main.py:156-160) -- harmless with an
    asynchronous parameter server, fatal for a synchronous exchange, where a rank that stops early leaves the others
    waiting in the collective.  Protocol: every ``window`` updates (the update count is the same on every rank, because
    updates are synchronous) the ranks all-reduce a "my loop has finished" bit over a CPU (gloo) group; a rank whose
    loop has finished keeps making updates from its replay shard until a check-in says that EVERY rank has finished.
    All ranks therefore leave after the same update."""

    def __init__(self, window=64, group=None):
        import torch.distributed as dist
        self.window = int(window)
        self.finished = False          # this rank's host loop is done
        self.all_finished = not is_distributed()
        self._group = group
        if is_distributed() and group is None:
            self._group = dist.new_group(backend='gloo') if dist.get_backend() != 'gloo' else dist.group.WORLD

    def after_update(self, update_count):
        """Call after every update with the number of updates made so far; returns True once every rank has finished."""
        if self.all_finished or update_count % self.window != 0:
            return self.all_finished
        import torch
        import torch.distributed as dist
        t = torch.tensor([1 if self.finished else 0], dtype=torch.int32)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self._group)
        self.all_finished = bool(int(t.item()))
        return self.all_finished
