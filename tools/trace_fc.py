import ctypes as C, os, sys, random
os.environ["DQN_S2D_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "dist-dqn_b200"))
import torch
import __graft_entry__ as ge
ge.build()
import bench, _cabi as c, network, replay_memory
dev = torch.device("cuda", 0); torch.cuda.set_device(0)
cfg = bench.atari_config(32, "cnn", "bf16"); cfg.device_index = 0
net = network.Network.create_network(cfg, (84, 84, 4), 6)
class A: capacity = 60000
rm = bench._make_ring(A, torch, c, replay_memory, 0, 0, dev)
for i in range(5):
    idx = torch.tensor(random.sample(range(rm.size()), 32), dtype=torch.int32, device=dev)
    net.train_step(rm, 32, d_indices=idx)
torch.cuda.synchronize()
n = 60 * 8 + 160 * 4 + 16
buf = (C.c_uint64 * n)()
c.lib.dqn_debug_trace.argtypes = [C.c_void_p, C.c_int32]
assert c.lib.dqn_debug_trace(buf, n) == 0
names = ["entry", "pdl passed", "x loaded", "fwd mma done", "epi done", "B1", "TD done", "B2", "dh done", "B3", "dh loaded", "dgrad mma done", "epi2 done", "exit"]
t0 = buf[1080 + 1]
for i, nm in enumerate(names):
    print("%-16s %8d clk" % (nm, buf[1080 + i] - t0))
