#!/usr/bin/env python3
"""bench.py -- DQN minibatch updates/sec on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py --gpus 1 --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's path on the host CPU (oracle port)
    torchrun --nproc-per-node N ... bench.py --gpus N ...    # data parallel, one rank per GPU

A "step" is one minibatch update: sample gather (replay ring -> 84x84x4 stacks) -> target + online
forward -> TD target -> loss -> backward -> optimizer, at the ATARI preset (scripts/dqn_params.sh:24-42).
`value`  : steps/s with the replay ring and the sample indices already in HBM (one CUDA-graph launch per step).
`e2e`    : steps/s through the public Python API (ReplayMemory.add x update_freq from host frames +
           DQNAgent._train_minibatch with host-side random.sample and a loss read-back every step).
`roofline`: the dominant kernel of the step, timed with CUDA events in an instrumented eager pass.
`cpu_baseline`: the oracle's transcription of the reference step (torch-CPU fp32, all host threads), rank 0, N=1.
"""
from __future__ import annotations

import argparse
import json
import os
import random
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "dist-dqn_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

A_PONG = 6
FRAME = 84 * 84
# per-sample MACs of the two conv architectures
MACS = {"cnn": [3612672, 1179648, 331776, 65536], "nature": [3276800, 2654208, 1806336, 1605632]}
FC = {"cnn": 256, "nature": 512}
# algorithmic bytes per sampled transition of the gather: write 2 stacks, read 5 unique frames, record+index+scalars
GATHER_BYTES = 2 * 4 * FRAME + 5 * FRAME + (32 + 4 + 13)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--arch", default="cnn", choices=["cnn", "nature"])
    ap.add_argument("--batch", type=int, default=None, help="minibatch per GPU (default 32; 512 = large-batch sweep)")
    ap.add_argument("--capacity", type=int, default=1000000, help="replay transitions per GPU")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--dp", default="peer", choices=["peer", "nccl"],
                    help="gradient exchange at N>1: fused into the optimizer over NVLink peer memory, or NCCL all-reduce")
    ap.add_argument("--precision", default="bf16", choices=["fp32", "bf16"],
                    help="bf16 = tcgen05 tensor-core mode (parity 2e-2, default; cnn arch); fp32 = CUDA-core parity mode (1e-4)")
    ap.add_argument("--cpu-steps", type=int, default=20)
    ap.add_argument("--e2e-steps", type=int, default=2000)
    ap.add_argument("--skip-cpu", action="store_true")
    return ap.parse_args()


def atari_config(batch, arch, precision="fp32"):
    return argparse.Namespace(
        precision=precision,
        network=arch, optimizer="rmsprop", lr=0.00025, reg_param=0.001, momentum=0.9, rmsprop_decay=0.95,
        reward_discount=0.99, minibatch_size=batch, max_batch=batch, loss="mse", seed=0, input_dtype="u8",
        device_index=None, sync=True, task_id=0, frames_per_state=4, update_freq=4, target_update_freq=10000,
        replay_start_size=0, resize_width=0, resize_height=0, init_random_action_prob=1.0, min_random_action_prob=0.1,
        random_action_explore_steps=1000000, summary_freq=1, logdir=tempfile.mkdtemp(prefix="dqn_bench_"),
        cuda_graph=True)


# ----------------------------------------------------------------------------------------------- clocks
class ClockSampler(object):
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.idx = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.QUERY,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except Exception:
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7),
                              ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------- ring fill
def fill_ring(rm_handle, capacity, seed, dev, stream):
    """Synthetic Atari-shaped replay: uniform random bytes, episodes of length
    U{50..2000}, one terminal per episode, rewards U{-1,0,1}, 6 actions.  Frames are produced on the device
    and appended with dqn_replay_append_frames_device."""
    import torch
    import _cabi as c
    rng = np.random.default_rng(seed)
    gen = torch.Generator(device=dev).manual_seed(seed)
    added = 0
    chunk_frames = 32768
    pos = c.c_i64()
    while added < capacity:
        # plan episodes for this chunk
        lens = []
        total = 0
        while total + 4 < chunk_frames and added + sum(lens) < capacity:
            L = int(rng.integers(50, 2001))
            L = min(L, capacity - added - sum(lens), chunk_frames - total - 4)
            if L <= 0:
                break
            lens.append(L)
            total += L + 4
        frames = torch.randint(0, 256, (total, FRAME), dtype=torch.uint8, device=dev, generator=gen)
        c.check(c.lib.dqn_replay_append_frames_device(rm_handle, frames.data_ptr(), total, c.C.byref(pos), stream))
        p = pos.value
        for L in lens:
            acts = rng.integers(0, A_PONG, L)
            rews = rng.integers(-1, 2, L)
            for i in range(L):
                c.lib.dqn_replay_add(rm_handle, p + 3 + i, p + 4 + i, int(acts[i]), float(rews[i]), 1 if i == L - 1 else 0)
            p += L + 4
            added += L
        c.check(c.lib.dqn_replay_flush(rm_handle, stream))
        torch.cuda.synchronize()
    return added


# ----------------------------------------------------------------------------------------------- reference arm
def reference_steps_per_s(arch, batch, steps, warmup, capacity=1000000, pool=8192, seed=0):
    """The reference's step on the host CPU: oracle port (TensorFlow cannot run on this image), all threads.
    deque of `capacity` tuples over a pool of `pool` distinct 84x84x4 uint8 stacks, 4 adds + 1 update per step."""
    import torch
    from oracle.replay_ref import ReplayRef
    from oracle.torch_ref import CpuBaselineAgent
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    torch.set_num_threads(cores)
    rng = np.random.default_rng(seed)
    stacks = [rng.integers(0, 256, (84, 84, 4), dtype=np.uint8) for _ in range(pool)]
    rm = ReplayRef(capacity)
    acts = rng.integers(0, A_PONG, capacity)
    rews = rng.integers(-1, 2, capacity)
    for i in range(capacity):
        rm.add(stacks[i % pool], int(acts[i]), float(rews[i]), stacks[(i + 1) % pool], (i % 997) == 996)
    agent = CpuBaselineAgent(arch, (84, 84, 4), A_PONG, "rmsprop", 0.00025, 0.001, 0.99, 0.95, 10000, seed=seed)
    random.seed(seed)
    j = capacity

    def one():
        nonlocal j
        for _ in range(4):
            rm.add(stacks[j % pool], int(acts[j % capacity]), float(rews[j % capacity]), stacks[(j + 1) % pool], False)
            j += 1
        return agent.train_minibatch(rm, batch)

    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = time.perf_counter() - t0
    return steps / dt, dt / steps * 1e3, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = args.batch or 32
    steps = args.steps or 20
    warmup = args.warmup if args.warmup is not None else 3
    v, ms, cores = reference_steps_per_s(args.arch, batch, steps, warmup)
    sample = "%d updates of batch %d (4 adds + 1 update each) on a 1M-entry deque over 8192 distinct stacks" % (steps, batch)
    line = {
        "impl": "reference", "metric": "Nature-DQN minibatch updates/sec", "value": v, "unit": "updates/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.arch, batch, args.gpus, 1000000),
        "cpu_baseline": {"value": v, "unit": "updates/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference-on-CPU (restated): TensorFlow is not installable here; oracle/torch_ref.py transcribes "
                "dqn_agent.py:108-143 on torch-CPU fp32",
    }
    emit(line)


def workload_config(arch, batch, n, capacity):
    return {"workload": "DQN minibatch update, %s conv net (%s), Pong shape: synthetic 84x84x4 uint8 stacks, 6 actions, "
                        "minibatch %d/GPU, %d-transition replay ring/GPU, RMSProp lr 2.5e-4 decay 0.95, gamma 0.99"
                        % ("reference" if arch == "cnn" else "Nature-paper", arch, batch, capacity),
            "arch": arch, "minibatch_per_gpu": batch, "global_batch": batch * n, "replay_capacity_per_gpu": capacity,
            "parallelism": "dp%d" % n,
            "cache": "inputs larger than L2: each step gathers fresh random transitions from a %.1f GB frame ring"
                     % (capacity * FRAME / 1e9)}


# ----------------------------------------------------------------------------------------------- own arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    ge.build()
    import _cabi as c
    import network
    import replay_memory
    from dqn_agent import DQNAgent
    from frame_buffer import FrameBuffer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus %d needs torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    batch = args.batch or 32            # same per-GPU minibatch at every N (weak scaling); --batch 512 = large-batch sweep
    steps = args.steps or (2000 if batch <= 64 else 300)
    warmup = max(3, args.warmup if args.warmup is not None else 50)
    if args.arch not in ("cnn", "nature"):
        args.precision = "fp32"                     # the tensor-core kernels cover the two conv nets
    cfg = atari_config(batch, args.arch, args.precision)
    cfg.device_index = local

    net = network.Network.create_network(cfg, (84, 84, 4), A_PONG, num_replicas=world)
    if world > 1:
        dist.broadcast(net.flat_buffer("online"), 0)
        net.sync_target()
        net.set_hyper(cfg, grad_scale=1.0 / world)
        if args.dp == "peer":
            net.connect_peers()
    rm = replay_memory.ReplayMemory(args.capacity, device=local, trusted=True)
    # create the ring through the shell (first add fixes shape/dtype), then bulk-fill it on the device
    seed_state = np.zeros((84, 84, 4), np.uint8)
    rm._ensure(seed_state)
    side = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(side):
        fill_ring(rm.handle, args.capacity, 100 + rank, dev, c.current_stream())
        torch.cuda.synchronize()
        rm._resync()
        size = rm.size()
        assert size == args.capacity, size

        # sample indices for every step, drawn with the reference's own sampler and uploaded before timing
        random.seed(1234 + rank)
        total_steps = warmup + steps
        idx_table = torch.tensor([random.sample(range(size), batch) for _ in range(total_steps)], dtype=torch.int32,
                                 device=dev)
        d_idx = torch.empty(batch, dtype=torch.int32, device=dev)
        grad = net.flat_buffer("grad")

        def eager_step(want_loss=False):
            if world == 1:
                net.train_step(rm, batch, d_indices=d_idx, want_loss=want_loss)
            elif net.peers_connected:
                net.dp_train_step(rm, batch, d_indices=d_idx, want_loss=want_loss)
            else:
                b = rm.gather(d_idx)
                net.backward_on_batch(b, want_loss=want_loss)
                dist.all_reduce(grad)
                net.apply_gradients()

        # kernels per step (eager launches are counted by the library)
        d_idx.copy_(idx_table[0])
        eager_step()
        torch.cuda.synchronize()
        n0 = c.lib.dqn_launch_count()
        eager_step()
        torch.cuda.synchronize()
        launches_per_step = c.lib.dqn_launch_count() - n0

        graph = None
        if not args.no_graph:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                eager_step()

        def step(i):
            d_idx.copy_(idx_table[i], non_blocking=True)
            if graph is not None:
                graph.replay()
            else:
                eager_step()

        for i in range(warmup):
            step(i)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(warmup, warmup + steps):
            step(i)
            if (i - warmup) % 10000 == 9999:
                net.sync_target()                       # ATARI preset: target sync every 10000 updates
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        clocks = sampler.stop() if rank == 0 else None
        ms_total = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms_total], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_total = float(t.item())
        ms_per_step = ms_total / steps
        # whole-job aggregate: per-GPU minibatches of `batch` samples processed per second, summed over ranks
        # (one synchronous update = `world` such minibatches)
        value = world * steps / (ms_total / 1e3)
        samples_per_s = world * batch * steps / (ms_total / 1e3)

        # ---- per-kernel durations (instrumented eager pass over fresh indices) -> roofline of the dominant kernel
        roof, kernels = None, {}
        if rank == 0 or world > 1:
            c.check(c.lib.dqn_net_set_timing(net.handle, 2))
            nprof = 30
            g_evt0, g_evt1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            gather_ms = 0.0
            for i in range(nprof):
                d_idx.copy_(idx_table[(warmup + i) % total_steps])
                if world > 1 and not net.peers_connected:
                    g_evt0.record()
                    b = rm.gather(d_idx)
                    g_evt1.record()
                    net.backward_on_batch(b)
                    dist.all_reduce(grad)
                    net.apply_gradients()
                else:
                    eager_step()
                c.check(c.lib.dqn_net_timing_collect(net.handle))
                if world > 1 and not net.peers_connected:
                    gather_ms += g_evt0.elapsed_time(g_evt1)
            c.check(c.lib.dqn_net_set_timing(net.handle, 0))
            tot, cnt = c.c_f64(), c.c_i32()
            s = 0
            while True:
                name = c.lib.dqn_stage_name(s)
                if not name:
                    break
                c.lib.dqn_net_timing_read(net.handle, s, c.C.byref(tot), c.C.byref(cnt))
                if cnt.value:
                    kernels[name.decode()] = tot.value / cnt.value * 1e3        # us per launch
                s += 1
            if world > 1 and not net.peers_connected:
                kernels["gather"] = gather_ms / nprof * 1e3
            roof = roofline(kernels, args.arch, batch, net.flat_size)

        # ---- the other precision mode, same workload, for the record (N=1 only; short)
        other = None
        if world == 1 and args.arch == "cnn" and not args.skip_cpu:
            ocfg = atari_config(batch, args.arch, "fp32" if args.precision == "bf16" else "bf16")
            ocfg.device_index = local
            onet = network.Network.create_network(ocfg, (84, 84, 4), A_PONG)
            d_idx.copy_(idx_table[0])
            for _ in range(3):
                onet.train_step(rm, batch, d_indices=d_idx)
            og = torch.cuda.CUDAGraph()
            with torch.cuda.graph(og, stream=side):
                onet.train_step(rm, batch, d_indices=d_idx)
            nother = min(steps, 500)
            for i in range(10):
                d_idx.copy_(idx_table[i], non_blocking=True); og.replay()
            o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            o0.record()
            for i in range(nother):
                d_idx.copy_(idx_table[warmup + i], non_blocking=True); og.replay()
            o1.record()
            torch.cuda.synchronize()
            other = {"precision": ocfg.precision, "value": nother / (o0.elapsed_time(o1) / 1e3), "unit": "updates/s",
                     "steps": nother}
            del og, onet

        # ---- gather-only throughput at a large batch (second half of BASELINE.json's metric)
        gather_info = None
        if rank == 0:
            gather_info = gather_sweep(rm, dev)

    # ---- end to end through the public API (host frames in, loss out, every step)
    e2e = run_e2e(args, cfg, net, rm, world, dev, DQNAgent, FrameBuffer, dist if world > 1 else None)

    # CUDA graphs that hold NCCL kernels must be gone before the communicator is torn down
    graph = None
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    if rank != 0:
        _finish(world, dist)
        return
    cpu = None
    if world == 1 and not args.skip_cpu:
        v, ms, cores = reference_steps_per_s(args.arch, batch, args.cpu_steps, 3)
        cpu = {"value": v, "unit": "updates/s", "cores": cores, "kind": "port",
               "sample": "%d updates of batch %d (4 adds + 1 update each) on a 1M-entry deque over 8192 distinct stacks; "
                         "oracle/torch_ref.py (torch-CPU fp32) because TensorFlow cannot run here" % (args.cpu_steps, batch)}
    line = {
        "metric": "Nature-DQN minibatch updates/sec", "value": value, "unit": "updates/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
        "config": workload_config(args.arch, batch, world, args.capacity),
        "samples_per_s": samples_per_s,
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches_per_step * steps),
        "launches_per_step": int(launches_per_step), "cuda_graph": not args.no_graph,
        "roofline": roof, "cpu_baseline": cpu, "kernels_us": kernels, "gather": gather_info,
        "precision": args.precision, "other_precision_mode": other,
        "dp_exchange": (args.dp if world > 1 else None),
    }
    emit(line)
    _finish(world, dist)




def _finish(world, dist):
    """Leave without running interpreter teardown: destroying NCCL communicators and CUDA graphs in arbitrary
    order at exit can hang."""
    sys.stdout.flush()
    sys.stderr.flush()
    if world > 1:
        try:
            dist.barrier()
        except Exception:
            pass
    os._exit(0)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


def _traffic(name, batch):
    """DRAM bytes per launch of `name` from the committed ncu --set full captures (taken at B=32 and B=512), else None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return t["dram_bytes_per_launch"].get(str(batch), {}).get(name)
    except Exception:
        return None


def roofline(kernels, arch, batch, nparams):
    """Roofline of the dominant kernel of the step (largest average duration)."""
    if not kernels:
        return None
    hbm, burst, sustained, src = peaks()
    # only kernels on the critical path of the step (the head / fc / conv3 / conv2 weight gradients and the loss run on
    # forked streams in the shadow of the data-gradient chain)
    side = ("head_wgrad", "loss", "wgrad_fc", "wgrad_conv3", "wgrad_conv2")
    crit = {k: v for k, v in kernels.items() if k not in side} or kernels
    name = max(crit, key=lambda k: crit[k])
    us = crit[name]
    macs = MACS[arch]
    layer = {"conv1": 0, "conv2": 1, "conv3": 2, "fc": 3}
    if name.startswith(("fwd_", "wgrad_", "dgrad_")):
        l = layer[name.split("_", 1)[1]]
        passes = 2 if name.startswith("fwd_") else 1          # online + target forward share one launch
        flops = 2.0 * macs[l] * batch * passes
        ach = flops / (us * 1e-6) / 1e12
        return {"kernel": name, "bound": "tensor", "achieved": ach, "peak": sustained, "unit": "TFLOP/s",
                "frac": ach / sustained, "traffic": _traffic(name, batch), "us_per_launch": us,
                "peak_source": src + " bf16 sustained",
                "note": "implicit GEMM measured against the bf16 tensor-pipe peak (fp32 mode runs on CUDA cores); durations "
                        "come from an instrumented eager pass (CUDA events around every launch, no PDL overlap), which adds "
                        "several microseconds per launch at batch 32 -- the fraction is meaningful at --batch 512"}
    if name == "gather":
        byts = GATHER_BYTES * batch
    elif name in ("optimizer", "reduce_grad"):
        byts = 20.0 * nparams
    else:
        byts = 0.0
    ach = byts / (us * 1e-6) / 1e9
    return {"kernel": name, "bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm,
            "traffic": None, "us_per_launch": us, "peak_source": src}


def gather_sweep(rm, dev, k=8192, iters=30):
    """Replay gather GB/s at a large batch (config 4 of BASELINE.json), CUDA-event timed, fresh indices each launch."""
    import torch
    import _cabi as c
    size = rm.size()
    random.seed(77)
    tables = torch.tensor([random.sample(range(size), k) for _ in range(iters + 3)], dtype=torch.int32, device=dev)
    states = torch.empty((k, 84, 84, 4), dtype=torch.uint8, device=dev)
    nexts = torch.empty_like(states)
    acts = torch.empty(k, dtype=torch.int32, device=dev)
    rews = torch.empty(k, dtype=torch.float64, device=dev)
    terms = torch.empty(k, dtype=torch.uint8, device=dev)
    st = c.current_stream()

    def go(i):
        c.check(c.lib.dqn_replay_gather(rm.handle, tables[i].data_ptr(), k, states.data_ptr(), nexts.data_ptr(),
                                        acts.data_ptr(), rews.data_ptr(), terms.data_ptr(), st))
    for i in range(3):
        go(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(3, 3 + iters):
        go(i)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    hbm, _, _, src = peaks()
    gbs = GATHER_BYTES * k / (us * 1e-6) / 1e9
    return {"k": k, "us_per_launch": us, "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / hbm, "peak_gbs": hbm,
            "bytes_per_transition": GATHER_BYTES, "peak_source": src}


def run_e2e(args, cfg, net, rm, world, dev, DQNAgent, FrameBuffer, dist):
    """updates/s through the public API with HOST inputs: per step update_freq x (FrameBuffer.append/get_state +
    ReplayMemory.add of a host frame) and DQNAgent._train_minibatch (host random.sample -> H2D indices -> CUDA graph
    -> loss read back to the host)."""
    import torch

    class _Env(object):            # DQNAgent only needs these at construction when replay_start_size == 0
        class spec:
            id = "SyntheticPong-v0"

        class action_space:
            n = A_PONG

            @staticmethod
            def sample():
                return 0

    agent = DQNAgent(_Env(), net, None, rm, cfg, enable_summary=True)
    batch = cfg.minibatch_size
    rng = np.random.default_rng(5)
    pool = rng.integers(0, 256, (256, 84, 84), dtype=np.uint8)
    fb = FrameBuffer(4)
    fb.append(pool[0])
    state = fb.get_state()
    n = args.e2e_steps
    warm = 10
    fi = 1

    def one():
        nonlocal state, fi
        for _ in range(cfg.update_freq):
            fb.append(pool[fi % 256]); fi += 1
            nxt = fb.get_state()
            rm.add(state, fi % A_PONG, float(fi % 3 - 1), nxt, False, newest=fb.frames[-1])     # as DQNAgent._record does
            state = nxt
        agent._train_minibatch(batch)          # summary_freq=1 -> the loss of every update is read back to the host
        return agent.last_loss                 # (asynchronously: it is picked up while the next update is prepared)

    for _ in range(warm):
        one()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(n):
        loss = one()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([dt], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    agent.flush_summaries()
    loss = agent.last_loss
    assert loss is not None and np.isfinite(loss)
    agent._graphs.clear()
    # the step graph copies a fixed prefix of the pinned staging buffer: 8 frame slots + header + 8 records, and the indices
    gi = rm.GRAPH_INGEST_FRAMES
    h2d = gi * FRAME + 64 + gi * 32 + batch * 4
    return {"value": world * n / dt, "unit": "updates/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
            "steps": n, "ms_per_step": dt / n * 1e3,
            "path": "FrameBuffer.append/get_state + ReplayMemory.add x%d (host uint8 frames) + DQNAgent._train_minibatch "
                    "(random.sample on host; one CUDA-graph launch per update moves the staged frames into the ring, "
                    "gathers, updates and copies the loss to pinned host memory, where it is collected one update later)"
                    % cfg.update_freq}


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else any library prints (NCCL's version banner
    goes to stdout) was redirected to stderr at start-up."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


if __name__ == "__main__":
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
