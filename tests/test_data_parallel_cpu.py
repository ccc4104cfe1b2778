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


# ---------------------------------------------------------------------------------------------------------------------
# Agent loop under data parallelism: ranks whose episodes have different lengths must still make the SAME number of
# gradient exchanges (a synchronous exchange hangs otherwise).  The network / replay memory are stand-ins; the exchange
# is a gloo all-reduce, which blocks for ever if one rank makes fewer calls than the other.
def _agent_worker(rank, world, port, out_q):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    import argparse
    import tempfile
    import torch.distributed as dist
    import data_parallel as dp
    import envs
    dp.init(backend="gloo")
    from dqn_agent import DQNAgent

    class FakeNet(object):
        peers_connected = False

        def sync_target(self):
            pass

        def predict(self, states, argmax=False):
            return None, torch.zeros(1, dtype=torch.int32)

    class FakeReplay(object):
        def __init__(self):
            self.n = 0

        def add(self, *a, **k):
            self.n += 1

        def size(self):
            return self.n

        def sample_indices(self, k):
            return list(range(k))

    class Agent(DQNAgent):
        exchanges = 0

        def _device_step(self, k, indices, want_loss):
            t = torch.ones(1)
            dist.all_reduce(t)                         # the gradient exchange: needs every rank
            assert t.item() == world
            self.exchanges += 1
            return t

    cfg = argparse.Namespace(
        minibatch_size=8, frames_per_state=1, resize_width=0, resize_height=0, update_freq=1, replay_start_size=0,
        init_random_action_prob=0.5, min_random_action_prob=0.1, random_action_explore_steps=100, target_update_freq=7,
        summary_freq=0, logdir=tempfile.mkdtemp(prefix="dqn_dp_agent_"), cuda_graph=False, dp_stop_window=5)
    # rank 0: short episodes, rank 1: long ones -> very different update counts per episode loop
    env = envs.CartPoleEnv(seed=rank, max_steps=6 if rank == 0 else 40)
    agent = Agent(env, FakeNet(), None, FakeReplay(), cfg, enable_summary=False)
    assert agent.replay_memory.size() >= cfg.minibatch_size          # the prefill covers one minibatch in data-parallel mode
    agent.train(num_episodes=4, max_steps_per_episode=50)
    counts = [None] * world
    dist.all_gather_object(counts, (agent.exchanges, agent.training_steps))
    if rank == 0:
        out_q.put(counts)
    dist.barrier()
    dist.destroy_process_group()


def test_agent_loops_of_unequal_length_make_the_same_number_of_exchanges():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_agent_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    counts = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert counts[0] == counts[1] and counts[0][0] == counts[0][1] > 0, counts
    assert counts[0][0] % 5 == 0                        # everyone left at a check-in (window of 5 updates)
