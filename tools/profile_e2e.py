"""Host-side profile of the end-to-end update loop (ReplayMemory.add x4 + DQNAgent._train_minibatch)."""
import cProfile
import os
import pstats
import random
import sys
import time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "dist-dqn_b200"))
import numpy as np
import torch
import __graft_entry__ as ge
ge.build()
import bench, _cabi as c, network, replay_memory
from dqn_agent import DQNAgent
dev = torch.device("cuda", 0); torch.cuda.set_device(0)
cfg = bench.atari_config(32, "cnn", "bf16"); cfg.device_index = 0
net = network.Network.create_network(cfg, (84, 84, 4), 6)


class A: capacity = 100000
rm = bench._make_ring(A, torch, c, replay_memory, 0, 0, dev)


class _Env(object):
    class spec: id = "x"
    class action_space:
        n = 6
        @staticmethod
        def sample(): return 0
agent = DQNAgent(_Env(), net, None, rm, cfg, enable_summary=True)
rng = np.random.default_rng(5)
NP = 256
pool = rng.integers(0, 256, (NP, 84, 84), dtype=np.uint8)
stacks = [np.ascontiguousarray(np.stack([pool[(i - 3 + c_) % NP] for c_ in range(4)], axis=-1)) for i in range(NP)]
fi = 0


def one():
    global fi
    for _ in range(4):
        rm.add(stacks[fi % NP], fi % 6, float(fi % 3 - 1), stacks[(fi + 1) % NP], False)
        fi += 1
    agent._train_minibatch(32)


for _ in range(50):
    one()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3000):
    one()
torch.cuda.synchronize()
print("us per update: %.1f" % ((time.perf_counter() - t0) / 3000 * 1e6))
# pieces
t0 = time.perf_counter()
for _ in range(3000):
    random.sample(range(100000), 32)
print("random.sample: %.1f us" % ((time.perf_counter() - t0) / 3000 * 1e6))
t0 = time.perf_counter()
for _ in range(3000):
    for _ in range(4):
        rm.add(stacks[fi % NP], fi % 6, float(fi % 3 - 1), stacks[(fi + 1) % NP], False)
        fi += 1
    rm.seal()
print("4 adds + seal: %.1f us" % ((time.perf_counter() - t0) / 3000 * 1e6))
pr = cProfile.Profile()
pr.enable()
for _ in range(3000):
    one()
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
# device timeline of three consecutive end-to-end updates (CUPTI through torch.profiler)
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(8):
        one()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
ing = [i for i, e in enumerate(evs) if "ingest" in e.name]
if len(ing) >= 5:
    t0 = evs[ing[3]].time_range.start
    print("device timeline of one end-to-end update (us from its ingest kernel): start, duration, name")
    for e in evs[ing[3]:ing[4] + 1]:
        print("%8.2f %7.2f  %s" % (e.time_range.start - t0, e.time_range.end - e.time_range.start, e.name[:80]))
    print("period: %s" % ["%.1f" % (evs[ing[i + 1]].time_range.start - evs[ing[i]].time_range.start) for i in range(len(ing) - 1)])
