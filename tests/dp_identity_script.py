"""Run under torchrun (one rank per GPU): an N-rank data-parallel update == a 1-rank update on the concatenated batch
(reference semantics: mean of the N replica gradients, one apply -- src/network.py:187-192), in both arithmetic modes,
through both exchanges (NCCL all-reduce; peer memory fused into the step's kernels, fed from each rank's own replay ring),
starting from DIFFERENT per-rank initial weights that the start-up broadcast must overwrite."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dist-dqn_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _ring_from_batch(replay_memory, batch, local):
    s, a, r, s2, t = batch
    rm = replay_memory.ReplayMemory(4 * len(a), device=local)
    for i in range(len(a)):
        rm.add(s[i], int(a[i]), float(r[i]), s2[i], bool(t[i]))
    return rm


def check_mode(precision, torch, dist, dp, network, replay_memory, helpers, rank, local, world):
    make_batch, make_config, oracle_params, rel_err, to_device_batch = helpers
    arch, ishape, A, B, steps = "cnn", (84, 84, 4), 6, 8, 3
    tol = 2e-5 if precision == "fp32" else 2e-3
    cfg = make_config(network=arch, minibatch_size=B * world, max_batch=B * world, device_index=local, precision=precision,
                      seed=100 + rank)                                 # every rank starts from different weights
    params = oracle_params(arch, ishape, A, 31)
    batches = [make_batch(ishape, A, B, 700 + 10 * i + rank) for i in range(steps)]
    gathered = [None] * world
    dist.all_gather_object(gathered, batches)

    def fresh(mode):
        net = network.Network.create_network(cfg, ishape, A, num_replicas=world)
        if rank == 0:
            net.set_params(params)
        dp.setup_network(net, cfg, mode=mode)                          # broadcast (+ bf16 operand refresh), 1/world, peers
        return net

    # single-rank reference on the concatenated batches (rank 0 only)
    ref_params = None
    if rank == 0:
        ref = network.Network.create_network(cfg, ishape, A)
        ref.set_params(params)
        ref.sync_target()
        for i in range(steps):
            cat = [np.concatenate([g[i][j] for g in gathered]) for j in range(5)]
            ref.train_on_batch(to_device_batch(ref, *cat))
        torch.cuda.synchronize()
        ref_params = ref.get_params()

    ok = True
    flats = {}
    for mode in ("nccl", "peer"):
        net = fresh(mode)
        rings = [_ring_from_batch(replay_memory, b, local) for b in batches]
        idx = torch.arange(B, dtype=torch.int32, device=net.device)
        for i in range(steps):
            dp.train_step(net, rings[i], B, idx)                       # ring -> gather -> update (dqn_train_step_dp for 'peer')
        torch.cuda.synchronize()
        if mode == "peer":
            net.dp_status()
        flat = net.flat_buffer("online").clone()
        flats[mode] = flat
        ref0 = flat.clone()
        dist.broadcast(ref0, 0)
        same = bool(torch.equal(flat, ref0))                           # bit-identical weights on every rank
        q = net.predict(gathered[0][0][0])                             # same probe states on every rank; runs on the refreshed bf16 operand images
        q0 = q.clone()
        dist.broadcast(q0, 0)
        same_q = bool(torch.equal(q, q0))
        err = 0.0
        if rank == 0:
            err = max(rel_err(a, b) for a, b in zip(net.get_params(), ref_params))
            print("%s %s: %d-rank x %d steps vs 1 rank on the concatenated batches: max rel err %.3g (tol %g); identical "
                  "weights %s, identical Q %s" % (precision, mode, world, steps, err, tol, same, same_q))
        ok = ok and same and same_q and err < tol
    perr = float(((flats["peer"] - flats["nccl"]).abs().max() / flats["nccl"].abs().max()).item())
    if rank == 0:
        print("%s: peer exchange vs NCCL exchange, max rel err %.3g" % (precision, perr))
    return ok and perr < tol


def main():
    import torch
    import torch.distributed as dist
    import data_parallel as dp
    import network
    import replay_memory
    from gpu_helpers import make_batch, make_config, oracle_params, rel_err, to_device_batch
    rank, local, world = dp.world_info()
    torch.cuda.set_device(local)
    dp.init(device=torch.device("cuda", local))
    helpers = (make_batch, make_config, oracle_params, rel_err, to_device_batch)
    ok = True
    for precision in ("fp32", "bf16"):
        ok = check_mode(precision, torch, dist, dp, network, replay_memory, helpers, rank, local, world) and ok
    t = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(t)
    dist.barrier()
    dist.destroy_process_group()
    if t.item() != 0:
        print("FAILED on some rank")
        sys.exit(1)
    if rank == 0:
        print("DP_IDENTITY_OK")


if __name__ == "__main__":
    main()
