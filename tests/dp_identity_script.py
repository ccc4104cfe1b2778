"""Run under torchrun (one rank per GPU): N-rank data-parallel step == 1-rank step on the concatenated batch."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dist-dqn_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    import torch
    import torch.distributed as dist
    import data_parallel as dp
    import network
    from gpu_helpers import make_batch, make_config, oracle_params, rel_err, to_device_batch
    rank, local, world = dp.world_info()
    torch.cuda.set_device(local)
    dp.init(device=torch.device("cuda", local))
    arch, ishape, A, B = "cnn", (84, 84, 4), 6, 8
    cfg = make_config(network=arch, minibatch_size=B * world, max_batch=B * world, device_index=local)
    net = network.Network.create_network(cfg, ishape, A, num_replicas=world)
    params = oracle_params(arch, ishape, A, 31)
    if rank == 0:
        net.set_params(params)
    dp.broadcast_parameters(net.flat_buffer("online"))
    net.sync_target()
    net.set_hyper(cfg, grad_scale=dp.grad_scale(world))
    batch = make_batch(ishape, A, B, 700 + rank)
    net.backward_on_batch(to_device_batch(net, *batch))
    dp.sync_gradients(net.flat_buffer("grad"))
    net.apply_gradients()
    torch.cuda.synchronize()
    got = net.get_params()
    gathered = [None] * world
    dist.all_gather_object(gathered, batch)
    ok = True
    if rank == 0:
        ref = network.Network.create_network(cfg, ishape, A)
        ref.set_params(params)
        ref.sync_target()
        cat = [np.concatenate([g[i] for g in gathered]) for i in range(5)]
        ref.train_on_batch(to_device_batch(ref, *cat))
        torch.cuda.synchronize()
        errs = [rel_err(a, b) for a, b in zip(got, ref.get_params())]
        print("dp identity max rel err", max(errs))
        ok = max(errs) < 2e-5
    # same step again through the peer-memory path (gradient exchange fused into the optimizer kernel)
    pnet = network.Network.create_network(cfg, ishape, A, num_replicas=world)
    if rank == 0:
        pnet.set_params(params)
    dp.broadcast_parameters(pnet.flat_buffer("online"))
    pnet.sync_target()
    pnet.connect_peers()
    for _ in range(3):                                   # several steps: exercises the double buffering / flags
        pnet.dp_train_on_batch(to_device_batch(pnet, *batch))
    torch.cuda.synchronize()
    for _ in range(2):
        net.backward_on_batch(to_device_batch(net, *batch))
        dp.sync_gradients(net.flat_buffer("grad"))
        net.apply_gradients()
    torch.cuda.synchronize()
    perr = max(rel_err(a, b) for a, b in zip(pnet.get_params(), net.get_params()))
    if rank == 0:
        print("peer path vs NCCL path after 3 steps, max rel err", perr)
    ok = ok and perr < 2e-5
    pflat = pnet.flat_buffer("online").clone()
    p0 = pflat.clone()
    dist.broadcast(p0, 0)
    ok = ok and bool(torch.equal(pflat, p0))
    # all ranks must hold identical weights afterwards
    flat = net.flat_buffer("online").clone()
    ref0 = flat.clone()
    dist.broadcast(ref0, 0)
    same = bool(torch.equal(flat, ref0))
    dist.barrier()
    dist.destroy_process_group()
    if not (ok and same):
        print("FAILED ok=%s same=%s" % (ok, same))
        sys.exit(1)
    if rank == 0:
        print("DP_IDENTITY_OK")


if __name__ == "__main__":
    main()
