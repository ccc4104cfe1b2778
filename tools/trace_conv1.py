"""Pipeline trace of the conv1 forward kernel (CTA 0): DQN_S2D_TRACE=1 python tools/trace_conv1.py [--batch 512]"""
import ctypes as C
import os
import sys
os.environ["DQN_S2D_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "dist-dqn_b200")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import argparse
import random
ap = argparse.ArgumentParser(); ap.add_argument("--batch", type=int, default=512); a = ap.parse_args()
import torch
import __graft_entry__ as ge
ge.build()
import bench, _cabi as c, network, replay_memory
dev = torch.device("cuda", 0); torch.cuda.set_device(0)
cfg = bench.atari_config(a.batch, "cnn", "bf16"); cfg.device_index = 0
net = network.Network.create_network(cfg, (84, 84, 4), 6)
a.capacity = 60000
rm = bench._make_ring(a, torch, c, replay_memory, 0, 0, dev)
for i in range(4):
    idx = torch.tensor(random.sample(range(rm.size()), a.batch), dtype=torch.int32, device=dev)
    net.train_step(rm, a.batch, d_indices=idx)
torch.cuda.synchronize()
n = 60 * 8 + 160 * 4
buf = (C.c_uint64 * n)()
c.lib.dqn_debug_trace.argtypes = [C.c_void_p, C.c_int32]
assert c.lib.dqn_debug_trace(buf, n) == 0
t0 = buf[0]
names = ["load issue", "raw_full", "img_empty", "build done", "mma start", "mma issued", "acc_full", "epi done"]
print("tile " + " ".join("%11s" % x for x in names))
for t in range(8):
    print("%4d " % t + " ".join("%11d" % (buf[t * 8 + e] - t0) for e in range(8)))

import numpy as np
ct = np.array([buf[480 + i] for i in range(148 * 4)], dtype=np.int64).reshape(148, 4)
ct = ct[ct[:, 0] > 0]
t0 = ct[:, 0].min()
print("CTAs %d: entry %d..%d ns, setup done +%d (median), pdl wait passed at %d..%d, exit %d..%d (median %d)" % (
    len(ct), 0, ct[:, 0].max() - t0, int(np.median(ct[:, 1] - ct[:, 0])), (ct[:, 2] - t0).min(), (ct[:, 2] - t0).max(),
    (ct[:, 3] - t0).min(), (ct[:, 3] - t0).max(), int(np.median(ct[:, 3] - t0))))
print("per-CTA busy (exit - pdl) ns: min %d median %d max %d" % ((ct[:, 3] - ct[:, 2]).min(), int(np.median(ct[:, 3] - ct[:, 2])), (ct[:, 3] - ct[:, 2]).max()))
# bracket (CUDA events around the launch, eager) of the same kernel, for comparison with the in-kernel span above
c.check(c.lib.dqn_net_set_timing(net.handle, 2))
for i in range(5):
    idx = torch.tensor(random.sample(range(rm.size()), a.batch), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    net.train_step(rm, a.batch, d_indices=idx)
    c.check(c.lib.dqn_net_timing_collect(net.handle))
    assert c.lib.dqn_debug_trace(buf, n) == 0
    ct = np.array([buf[480 + j] for j in range(148 * 4)], dtype=np.int64).reshape(148, 4)
    print("step %d: in-kernel span (first entry -> last exit) %.2f us" % (i, (ct[:, 3].max() - ct[:, 0].min()) / 1e3))
tot, cnt = c.c_f64(), c.c_i32()
c.lib.dqn_net_timing_read(net.handle, 1, C.byref(tot), C.byref(cnt))
print("eager event bracket fwd_conv1: %.2f us (mean of %d)" % (tot.value / cnt.value * 1e3, cnt.value))
fl = c.c_f64(); c.lib.dqn_net_timing_floor(net.handle, c.current_stream(), C.byref(fl)); print("empty-launch floor %.2f us" % fl.value)
# weight-gradient kernel, CTA 0 (stamps at 1136 + 8*tile + ev)
n2 = 60 * 8 + 160 * 4 + 16 + 64
buf2 = (C.c_uint64 * n2)()
assert c.lib.dqn_debug_trace(buf2, n2) == 0
wn = ["stack landed", "prev MMAs done", "image built", "grad landed", "dZ routed", "mma start", "mma issued", "(t0: MMAs done, t1: exit)"]
w0 = buf2[1136]
print("weight gradient, CTA 0 (clk)\ntile " + " ".join("%14s" % x for x in wn))
for t in range(5):
    print("%4d " % t + " ".join("%14d" % (buf2[1136 + t * 8 + e] - w0 if buf2[1136 + t * 8 + e] else -1) for e in range(8)))
