"""CPU, world_size 2 over gloo: the host-side data-parallel logic (gradient sum + 1/world scale + L2 once)
reproduces a single-rank step on the concatenated batch.  Compute here is the oracle (test-only use)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from conftest import PKG, ROOT


def _worker(rank, world, port, out_q):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    import data_parallel as dp
    from oracle import net_ref as N
    assert dp.init(backend="gloo") == world and dp.world_info() == (rank, rank, world)
    arch, ishape, A, B = "simple", (4,), 2, 6
    params = N.init_params(arch, ishape, A, seed=3, stddev=0.3)
    flat_p = torch.from_numpy(np.concatenate([p.ravel() for p in params]))
    if rank != 0:
        flat_p.zero_()
    dp.broadcast_parameters(flat_p)                      # rank 0 initialises, others copy
    assert np.array_equal(flat_p.numpy(), np.concatenate([p.ravel() for p in params]))
    rng = np.random.default_rng(50 + rank)
    s, y, a = rng.standard_normal((B, 4)), rng.standard_normal(B), rng.integers(0, A, B)
    _, _, g = N.loss_and_grads(arch, params, s, y, a, 0.0, ishape)            # local mean gradient, data term only
    flat_g = torch.from_numpy(np.concatenate([x.ravel() for x in g]))
    dp.sync_gradients(flat_g)                            # sum over ranks
    flat_g *= dp.grad_scale(world)
    gathered = [None] * world
    dist.all_gather_object(gathered, (s, y, a))
    if rank == 0:
        out_q.put((flat_g.numpy().copy(), gathered))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gradient_mean_equals_one_rank_on_concatenated_batch():
    from oracle import net_ref as N
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    flat_g, gathered = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    arch, ishape, A = "simple", (4,), 2
    params = N.init_params(arch, ishape, A, seed=3, stddev=0.3)
    s = np.concatenate([g[0] for g in gathered]); y = np.concatenate([g[1] for g in gathered])
    a = np.concatenate([g[2] for g in gathered])
    _, _, g_ref = N.loss_and_grads(arch, params, s, y, a, 0.0, ishape)
    ref = np.concatenate([x.ravel() for x in g_ref])
    assert np.allclose(flat_g, ref, rtol=1e-12, atol=1e-14)


def test_shard_seed_and_scale():
    import data_parallel as dp
    assert dp.grad_scale(8) == 0.125 and dp.shard_seed(None, 3) is None
    assert len({dp.shard_seed(7, r) for r in range(8)}) == 8
