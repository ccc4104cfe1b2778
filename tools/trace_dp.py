"""In-kernel globaltimer stamps of the data-parallel gradient exchange (peer-memory mode), one line per rank.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/trace_dp.py"""
import ctypes as C, os, sys, random
os.environ["DQN_S2D_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "dist-dqn_b200"))
import torch
import torch.distributed as dist
import __graft_entry__ as ge
ge.build()
import bench, _cabi as c, network, replay_memory
import data_parallel as dp
rank, local, world = dp.world_info()
dev = torch.device("cuda", local); torch.cuda.set_device(local)
dp.init(device=dev)
B = 32
cfg = bench.atari_config(B, "cnn", "bf16"); cfg.device_index = local
net = network.Network.create_network(cfg, (84, 84, 4), 6)
dp.setup_network(net, cfg, "peer")
class A: capacity = 60000
rm = bench._make_ring(A, torch, c, replay_memory, rank, local, dev)
random.seed(100 + rank)
tab = torch.tensor([random.sample(range(rm.size()), B) for _ in range(64)], dtype=torch.int32, device=dev)
d_idx = torch.empty(B, dtype=torch.int32, device=dev); d_idx.copy_(tab[0])
side = torch.cuda.Stream()
n = 60 * 8 + 160 * 4 + 16 + 64
buf = (C.c_uint64 * n)()
c.lib.dqn_debug_trace.argtypes = [C.c_void_p, C.c_int32]
names = {0: "conv3+head reduce: start", 1: "conv3+head reduce: last block", 2: "conv3+head reduce: flags raised",
         32: "conv3+head peer sum: start", 33: "conv3+head peer sum: flags seen", 34: "conv3+head peer sum: end",
         16: "fc reduce: start", 17: "fc reduce: last block", 18: "fc reduce: flags raised",
         40: "fc peer sum: start", 41: "fc peer sum: flags seen", 42: "fc peer sum: end",
         8: "conv1+conv2 reduce: start", 9: "conv1+conv2 reduce: last block", 10: "conv1+conv2 reduce: flags raised",
         48: "optimizer: start", 49: "optimizer: flags seen", 50: "optimizer: end"}
with torch.cuda.stream(side):
    for _ in range(3):
        dp.train_step(net, rm, B, d_idx)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        dp.train_step(net, rm, B, d_idx)
    rows = []
    for it in range(40):
        d_idx.copy_(tab[it % 64], non_blocking=True)
        g.replay()
        if it >= 30:
            torch.cuda.synchronize()
            assert c.lib.dqn_debug_trace(buf, n) == 0
            rows.append([buf[1136 + k] for k in sorted(names)])
torch.cuda.synchronize()
dist.barrier()
for r in range(world):
    if r == rank:
        print("rank %d (us relative to the start of the conv1+conv2 reduce = end of the backward pass, median of 10 updates)" % rank)
        keys = sorted(names)
        ref = keys.index(8)
        import statistics
        for j, k in enumerate(keys):
            print("  %-34s %8.2f" % (names[k], statistics.median((row[j] - row[ref]) / 1e3 for row in rows)))
        sys.stdout.flush()
    dist.barrier()
dist.destroy_process_group()
