"""Device timeline of one graph-replayed update (kernel start / duration from CUPTI through torch.profiler).
    python tools/timeline.py [--batch 32]
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/timeline.py
        (data-parallel step with the peer-memory exchange; rank 0 prints its timeline)"""
import argparse
import os
import random
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "dist-dqn_b200"))
ap = argparse.ArgumentParser(); ap.add_argument("--batch", type=int, default=32); ap.add_argument("--eager", action="store_true"); ap.add_argument("--copy-indices", action="store_true")
a = ap.parse_args()
import torch
import __graft_entry__ as ge
ge.build()
import bench, _cabi as c, network, replay_memory
import data_parallel as dp
from torch.profiler import profile, ProfilerActivity
rank, local, world = dp.world_info()
dev = torch.device("cuda", local); torch.cuda.set_device(local)
dp.init(device=dev)
cfg = bench.atari_config(a.batch, "cnn", "bf16"); cfg.device_index = local
net = network.Network.create_network(cfg, (84, 84, 4), 6)
dp.setup_network(net, cfg, "peer")
a.capacity = 100000
rm = bench._make_ring(a, torch, c, replay_memory, rank, local, dev)
random.seed(100 + rank)
tab = torch.tensor([random.sample(range(rm.size()), a.batch) for _ in range(64)], dtype=torch.int32, device=dev)
d_idx = torch.empty(a.batch, dtype=torch.int32, device=dev)
d_idx.copy_(tab[0])
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    for _ in range(3):
        dp.train_step(net, rm, a.batch, d_idx)
    cursor = replay_memory.IndexTable(tab, dev)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        if a.copy_indices:
            dp.train_step(net, rm, a.batch, d_idx)
        else:
            dp.train_step_table(net, rm, cursor)      # as bench.py: the update takes its own row of the index table

    def step(i):
        if a.eager or a.copy_indices:
            d_idx.copy_(tab[i % 64], non_blocking=True)
        if a.eager:
            dp.train_step(net, rm, a.batch, d_idx)
        else:
            g.replay()
    for i in range(20):
        step(i)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for i in range(6):
            step(i)
        torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
# split into steps at the index copy (memcpy DtoD) that precedes every update
steps, cur = [], []
for e in evs:
    if "Memcpy" in e.name or "memcpy" in e.name or "index_table_next" in e.name or (not a.copy_indices and "conv1_s2d_fwd" in e.name):
        if cur:
            steps.append(cur)
        cur = []
    cur.append(e)
steps.append(cur)
if world > 1:
    import torch.distributed as dist
    dist.barrier()
if rank != 0:
    dist.destroy_process_group()
    sys.exit(0)
st = steps[-2]
t0 = st[0].time_range.start
print("kernels of one update (us from the index copy / index kernel): start, duration, name")
for e in st:
    print("%8.2f %7.2f  %s" % (e.time_range.start - t0, e.time_range.end - e.time_range.start, e.name[:90]))
print("update period: %.2f us" % (steps[-1][0].time_range.start - t0))
if world > 1:
    dist.destroy_process_group()
