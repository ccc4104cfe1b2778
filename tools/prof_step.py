"""Small driver for ncu captures: a few sample-to-update steps at a given minibatch on a modest ring.
    python tools/prof_step.py --batch 512 --steps 5 [--precision bf16] [--arch cnn]"""
import argparse
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dist-dqn_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--capacity", type=int, default=60000)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--arch", default="cnn")
    ap.add_argument("--gather", type=int, default=0, help="also run this many plain gathers of --batch transitions")
    a = ap.parse_args()
    import numpy as np
    import torch
    import __graft_entry__ as ge
    ge.build()
    import bench
    import _cabi as c
    import network
    import replay_memory
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    cfg = bench.atari_config(a.batch, a.arch, a.precision)
    cfg.device_index = 0
    net = network.Network.create_network(cfg, (84, 84, 4), 6)
    a.capacity = max(a.capacity, 2 * a.batch)
    rm = bench._make_ring(a, torch, c, replay_memory, 0, 0, dev)
    random.seed(0)
    for i in range(a.steps):
        idx = torch.tensor(random.sample(range(rm.size()), a.batch), dtype=torch.int32, device=dev)
        net.train_step(rm, a.batch, d_indices=idx)
    for i in range(a.gather):
        idx = torch.tensor(random.sample(range(rm.size()), a.batch), dtype=torch.int32, device=dev)
        rm.gather(idx)
    torch.cuda.synchronize()
    print("done")


if __name__ == "__main__":
    main()
