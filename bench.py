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
    ap.add_argument("--config", type=int, default=2, choices=[1, 2, 3, 4, 5],
                    help="BASELINE.json config: 1 CartPole MLP (B=32 and 100), 2 conv net B=32 (default, the driver's line), "
                         "3 conv net B=512/GPU, 4 replay gather sweep k=32..8192, 5 Breakout shape: 1024-env greedy act + "
                         "B=512 step + target sync")
    ap.add_argument("--cpu-steps", type=int, default=100)
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
def reference_steps_per_s(arch, batch, steps, warmup, capacity=1000000, pool=8192, seed=0, one_thread_steps=0):
    """The reference's step on the host CPU: oracle port (TensorFlow cannot run on this image), all threads.
    deque of `capacity` tuples over a pool of `pool` distinct 84x84x4 uint8 stacks, 4 adds + 1 update per step.
    one_thread_steps > 0: afterwards the same loop with torch limited to one thread (returned as a 4th value)."""
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
    if one_thread_steps:
        torch.set_num_threads(1)
        one()
        t1 = time.perf_counter()
        for _ in range(one_thread_steps):
            one()
        single = one_thread_steps / (time.perf_counter() - t1)
        torch.set_num_threads(cores)
        return steps / dt, dt / steps * 1e3, cores, single
    return steps / dt, dt / steps * 1e3, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = args.batch or (512 if args.config in (3, 5) else 32)
    steps = args.steps or 20
    warmup = args.warmup if args.warmup is not None else 3
    # one timed "step" of this arm is a bounded sample of `per` updates, so that even --steps 20 times >= 100 updates
    per = max(1, -(-100 // steps)) if batch <= 64 else 1
    v, ms, cores = reference_steps_per_s(args.arch, batch, steps * per, warmup)
    sample = ("%d steps x %d updates of batch %d (4 adds + 1 update each) on a 1M-entry deque over 8192 distinct stacks"
              % (steps, per, batch))
    line = {
        "impl": "reference", "metric": "Nature-DQN minibatch updates/sec", "value": v, "unit": "updates/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": ms * per, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.arch, batch, args.gpus, 1000000),
        "cpu_baseline": {"value": v, "unit": "updates/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference-on-CPU (restated): TensorFlow is not installable here; oracle/torch_ref.py transcribes "
                "dqn_agent.py:108-143 on torch-CPU fp32; run-to-run spread on this pool's hosts is about 160-185 updates/s",
    }
    emit(line)


def workload_config(arch, batch, n, capacity):
    return {"workload": "DQN minibatch update, %s conv net (%s), Pong shape: synthetic 84x84x4 uint8 stacks, 6 actions, "
                        "minibatch %d/GPU, %d-transition replay ring/GPU, RMSProp lr 2.5e-4 decay 0.95, gamma 0.99"
                        % ("reference" if arch == "cnn" else "Nature-paper", arch, batch, capacity),
            "arch": arch, "minibatch_per_gpu": batch, "global_batch": batch * n, "replay_capacity_per_gpu": capacity,
            "parallelism": "dp%d" % n,
            "cache": "inputs larger than L2: each step gathers fresh random transitions from a %.1f GB frame ring"
                     % (capacity * FRAME / 1e9),
            "indices": "random.sample per step, drawn before timing and resident in HBM as one table; the captured "
                       "update reads its own row and advances the cursor (dqn_train_step_table) -- the host only "
                       "replays the graph"}


# ----------------------------------------------------------------------------------------------- own arm
def _setup(args):
    """Process-level set-up shared by every config: build, device, process group."""
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    ge.build()
    import data_parallel as dp
    rank, local, world = dp.world_info()
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("--gpus %d needs torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dp.init(device=dev)
    return torch, dist, dp, rank, local, world, dev


def _teardown(world, dist, torch):
    """Ordered shutdown (graphs that hold NCCL kernels and the peer mappings go before the communicator), then a
    normal return: the interpreter exits through its usual path."""
    import gc
    gc.collect()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sys.stdout.flush()
    sys.stderr.flush()


def _timed_steps(torch, dist, world, rank, local, dev, step, warmup, steps, after_step=None):
    """`warmup` untimed steps, then EXACTLY `steps` timed ones between CUDA events, bracketed by barrier + synchronize.
    The clock sampler (an nvidia-smi child of rank 0) is started BEFORE the pre-timing barrier, and three more untimed
    synchronised steps follow the barrier, so that no rank enters the timed region while another is still busy on the
    host.  Returns (ms_total as max over ranks, [per-rank ms], clocks)."""
    sampler = ClockSampler(local)
    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    if rank == 0:
        sampler.start()
    if world > 1:
        dist.barrier()
    for i in range(3):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(warmup, warmup + steps):
        step(i)
        if after_step is not None:
            after_step(i - warmup)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    ms = e0.elapsed_time(e1)
    per_rank = [ms]
    if world > 1:
        t = torch.zeros(world, device=dev)
        t[rank] = ms
        dist.all_reduce(t)
        per_rank = [float(x) for x in t.tolist()]
    return max(per_rank), per_rank, clocks


def _kernel_times(torch, c, net, run_one, nprof=30, stream=None, prepare=None):
    """us per launch of every kernel of a step: the step is captured into a CUDA graph with a CUDA-event record node on
    either side of every kernel, and the graph is replayed `nprof` times on fresh inputs (`prepare(i)`).  The events
    stamp each kernel as it runs inside the graph, so a bracket holds neither host launch latency nor the neighbours'
    overlap.  (stream=None: eager launches bracketed by ordinary event records -- several microseconds of launch
    latency per bracket.)"""
    c.check(c.lib.dqn_net_set_timing(net.handle, 2))
    if stream is not None:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            run_one(0)
            c.check(c.lib.dqn_net_timing_floor_enqueue(net.handle, c.current_stream()))
        c.check(c.lib.dqn_net_set_timing(net.handle, 3))
        for i in range(nprof):
            if prepare is not None:
                prepare(i)
            g.replay()
            c.check(c.lib.dqn_net_timing_collect(net.handle))
        del g
    else:
        for i in range(nprof):
            if prepare is not None:
                prepare(i)
            run_one(i)
            c.check(c.lib.dqn_net_timing_collect(net.handle))
    c.check(c.lib.dqn_net_set_timing(net.handle, 0))
    kernels = {}
    tot, cnt = c.c_f64(), c.c_i32()
    s = 0
    while True:
        name = c.lib.dqn_stage_name(s)
        if not name:
            break
        c.lib.dqn_net_timing_read(net.handle, s, c.C.byref(tot), c.C.byref(cnt))
        if cnt.value:
            kernels[name.decode()] = tot.value / cnt.value * 1e3
        s += 1
    if "launch_floor" in kernels:                     # in-graph brackets: the floor was measured inside the same graph
        return kernels, kernels.pop("launch_floor")
    floor = c.c_f64()
    c.check(c.lib.dqn_net_timing_floor(net.handle, c.current_stream(), c.C.byref(floor)))
    return kernels, floor.value


def _dp_check(torch, dist, dp, network, cfg, rm, batch, d_idx, precision):
    """Untimed self-check of the data-parallel step (the driver's test box has one GPU): from the same weights, one
    update through the peer-memory exchange and one through backward -> NCCL all-reduce -> apply must give the same
    weights, and every rank must hold bit-identical weights and bf16 operand images afterwards."""
    a = network.Network.create_network(cfg, (84, 84, 4), A_PONG, num_replicas=dist.get_world_size())
    b = network.Network.create_network(cfg, (84, 84, 4), A_PONG, num_replicas=dist.get_world_size())
    dp.setup_network(a, cfg, mode="peer")
    dp.setup_network(b, cfg, mode="nccl")
    dp.train_step(a, rm, batch, d_idx)
    dp.train_step(b, rm, batch, d_idx)
    torch.cuda.synchronize()
    wa, wb = a.flat_buffer("online").clone(), b.flat_buffer("online").clone()
    rel = float(((wa - wb).abs().max() / wb.abs().max()).item())
    ref = wa.clone()
    dist.broadcast(ref, 0)
    same = bool(torch.equal(ref, wa))
    probe = rm.gather(d_idx).states[:8].clone()           # the same 8 states on every rank (each rank's ring differs)
    dist.broadcast(probe, 0)
    q = a.predict(probe)                                  # through the refreshed bf16 operand images
    qr = q.clone()
    dist.broadcast(qr, 0)
    t = torch.tensor([rel, 0.0 if same else 1.0, 0.0 if torch.equal(q, qr) else 1.0], device=wa.device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    a.dp_status()
    ok = t[0].item() < 2e-5 and t[1].item() == 0.0 and t[2].item() == 0.0
    del a, b
    return {"result": "ok" if ok else "FAILED", "peer_vs_nccl_max_rel": float(t[0].item()),
            "weights_identical_across_ranks": t[1].item() == 0.0, "q_identical_across_ranks": t[2].item() == 0.0,
            "precision": precision}


def run_b200(args):
    torch, dist, dp, rank, local, world, dev = _setup(args)
    if args.config == 1:
        line = run_config1(args, torch, dev)
    elif args.config == 4:
        line = run_config4(args, torch, dev)
    elif args.config == 5:
        line = run_config5(args, torch, dist, dp, rank, local, world, dev)
    else:
        line = run_update_bench(args, torch, dist, dp, rank, local, world, dev)
    if rank == 0 and line is not None:
        emit(line)
    _teardown(world, dist, torch)


def _make_ring(args, torch, c, replay_memory, rank, local, dev):
    rm = replay_memory.ReplayMemory(args.capacity, device=local, trusted=True)
    rm._ensure(np.zeros((84, 84, 4), np.uint8))       # the first add fixes shape/dtype; then bulk-fill on the device
    fill_ring(rm.handle, args.capacity, 100 + rank, dev, c.current_stream())
    torch.cuda.synchronize()
    rm._resync()
    assert rm.size() == args.capacity, rm.size()
    return rm


def run_update_bench(args, torch, dist, dp, rank, local, world, dev):
    """Configs 2 (B=32, the driver's line) and 3 (B=512 per GPU)."""
    import _cabi as c
    import network
    import replay_memory
    from dqn_agent import DQNAgent
    from frame_buffer import FrameBuffer
    batch = args.batch or (512 if args.config == 3 else 32)      # same per-GPU minibatch at every N (weak scaling)
    steps = args.steps or (2000 if batch <= 64 else 300)
    warmup = max(3, args.warmup if args.warmup is not None else 50)
    if args.arch not in ("cnn", "nature"):
        args.precision = "fp32"                     # the tensor-core kernels cover the two conv nets
    cfg = atari_config(batch, args.arch, args.precision)
    cfg.device_index = local

    net = network.Network.create_network(cfg, (84, 84, 4), A_PONG, num_replicas=world)
    dp.setup_network(net, cfg, mode=args.dp)
    side = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(side):
        rm = _make_ring(args, torch, c, replay_memory, rank, local, dev)
        size = rm.size()
        # sample indices for every step, drawn with the reference's own sampler and uploaded before timing
        random.seed(1234 + rank)
        total_steps = warmup + 3 + steps          # warm-up, the 3 untimed synchronised steps of _timed_steps, timed steps: no row is used twice
        idx_table = torch.tensor([random.sample(range(size), batch) for _ in range(total_steps)], dtype=torch.int32,
                                 device=dev)
        d_idx = torch.empty(batch, dtype=torch.int32, device=dev)

        def eager_step(want_loss=False):
            dp.train_step(net, rm, batch, d_idx, want_loss=want_loss)

        dp_check = None
        if world > 1 and args.dp == "peer":
            d_idx.copy_(idx_table[0])
            dp_check = _dp_check(torch, dist, dp, network, cfg, rm, batch, d_idx, args.precision)

        # kernels per step (eager launches are counted by the library)
        d_idx.copy_(idx_table[0])
        eager_step()
        torch.cuda.synchronize()
        n0 = c.lib.dqn_launch_count()
        eager_step()
        torch.cuda.synchronize()
        launches_per_step = c.lib.dqn_launch_count() - n0

        # The captured update reads the step's row of the index table itself (fresh indices every step, the ring is far
        # larger than L2) and advances the table's device-side cursor: the host only replays the graph.
        graph = None
        copy_indices = os.environ.get("DQN_BENCH_COPY_INDICES") == "1"      # A/B: a D2D copy in front of every replay
        if not args.no_graph:
            cursor_table = replay_memory.IndexTable(idx_table, dev)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                if copy_indices:
                    eager_step()
                else:
                    dp.train_step_table(net, rm, cursor_table)      # the conv1 kernels read the row, the optimizer advances the cursor
            cursor_table.cursor.zero_()

        def step(i):
            if graph is not None and not copy_indices:
                graph.replay()
            else:
                d_idx.copy_(idx_table[i], non_blocking=True)
                graph.replay() if graph is not None else eager_step()

        def after(j):
            if j % 10000 == 9999:
                net.sync_target()                       # ATARI preset: target sync every 10000 updates

        ms_total, per_rank_ms, clocks = _timed_steps(torch, dist, world, rank, local, dev, step, warmup, steps, after)
        ms_per_step = ms_total / steps
        if world > 1 and dp_check is not None:
            # after the timed updates: every rank must still hold bit-identical weights (a stale read in the exchange
            # would have made the replicas drift apart for good), and no exchange wait may have timed out
            wa = net.flat_buffer("online").clone()
            ref = wa.clone()
            dist.broadcast(ref, 0)
            t = torch.tensor([0.0 if torch.equal(ref, wa) else 1.0], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            net.dp_status()
            dp_check["weights_identical_after_timed_steps"] = t[0].item() == 0.0
            if t[0].item() != 0.0:
                dp_check["result"] = "FAILED"
        # whole-job aggregate: per-GPU minibatches of `batch` samples processed per second, summed over ranks
        # (one synchronous update = `world` such minibatches)
        value = world * steps / (ms_total / 1e3)
        samples_per_s = world * batch * steps / (ms_total / 1e3)

        # ---- per-kernel durations (instrumented eager pass over fresh indices) -> roofline of the dominant kernel
        kernels, floor_us = _kernel_times(torch, c, net, lambda i: eager_step(), stream=None if args.no_graph else side,
                                          prepare=lambda i: d_idx.copy_(idx_table[(warmup + i) % total_steps]))
        roof = roofline(kernels, args.arch, batch, net.flat_size, floor_us)

        # ---- the same at the large batch (config 3's per-GPU batch), where the GEMM kernels are no longer latency-bound
        roof_large, large = None, None
        if world == 1 and batch != 512 and args.arch == "cnn" and not args.skip_cpu:
            roof_large, large = large_batch_block(args, torch, c, network, rm, dev, side)

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

        # ---- gather-only throughput (second half of BASELINE.json's metric)
        gather_info = None
        if rank == 0:
            gather_info = gather_sweep(rm, dev, ks=(512, 8192))

    # ---- end to end through the public API (host frames in, loss out, every step)
    e2e = run_e2e(args, cfg, net, rm, world, dev, DQNAgent, FrameBuffer, dist if world > 1 else None)
    if world > 1 and net.peers_connected:
        net.dp_status()

    # CUDA graphs that hold NCCL kernels must be gone before the communicator is torn down
    graph = None
    torch.cuda.synchronize()
    if rank != 0:
        return None
    cpu = None
    if world == 1 and not args.skip_cpu:
        v, ms, cores, single = reference_steps_per_s(args.arch, batch, args.cpu_steps if batch <= 64 else 10, 3,
                                                     one_thread_steps=20 if batch <= 64 else 3)
        cpu = {"value": v, "unit": "updates/s", "cores": cores, "kind": "port", "one_thread_value": single,
               "sample": "%d updates of batch %d (4 adds + 1 update each) on a 1M-entry deque over 8192 distinct stacks; "
                         "oracle/torch_ref.py (torch-CPU fp32) because TensorFlow cannot run here; run-to-run spread on "
                         "this pool's hosts is about 160-185 updates/s at batch 32" % (args.cpu_steps if batch <= 64 else 10, batch)}
    return {
        "metric": "Nature-DQN minibatch updates/sec", "value": value, "unit": "updates/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
        "config": workload_config(args.arch, batch, world, args.capacity),
        "samples_per_s": samples_per_s, "per_rank_ms_per_step": [m / steps for m in per_rank_ms],
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches_per_step * steps),
        "launches_per_step": int(launches_per_step), "cuda_graph": not args.no_graph,
        "roofline": roof, "roofline_large_batch": roof_large, "large_batch": large,
        "cpu_baseline": cpu, "kernels_us": kernels, "event_floor_us": floor_us,
        "gather": gather_info, "precision": args.precision, "other_precision_mode": other,
        "dp_exchange": (args.dp if world > 1 else None), "dp_check": dp_check,
    }


def large_batch_block(args, torch, c, network, rm, dev, side, batch=512, steps=100):
    """B=512 on the same ring in the same run: graph-replayed updates/s and the roofline of its dominant kernel."""
    cfg = atari_config(batch, args.arch, args.precision)
    cfg.device_index = dev.index
    net = network.Network.create_network(cfg, (84, 84, 4), A_PONG)
    size = rm.size()
    random.seed(4321)
    tab = torch.tensor([random.sample(range(size), batch) for _ in range(steps + 13)], dtype=torch.int32, device=dev)
    d_idx = torch.empty(batch, dtype=torch.int32, device=dev)
    d_idx.copy_(tab[0])
    for _ in range(3):
        net.train_step(rm, batch, d_indices=d_idx)
    import replay_memory
    cursor_table = replay_memory.IndexTable(tab, dev)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        net.train_step_table(rm, cursor_table)         # as in run_update_bench: the update takes its own row of indices
    cursor_table.cursor.zero_()
    for i in range(10):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps

    kernels, floor_us = _kernel_times(torch, c, net, lambda i: net.train_step(rm, batch, d_indices=d_idx), nprof=20, stream=side,
                                      prepare=lambda i: d_idx.copy_(tab[i % (steps + 13)]))
    roof = roofline(kernels, args.arch, batch, net.flat_size, floor_us)
    flops = 2.0 * batch * (4 * sum(MACS[args.arch]) - MACS[args.arch][0])    # fwd x2 + wgrad + dgrad (no conv1 dgrad)
    del g, net
    return roof, {"minibatch": batch, "value": 1e3 / ms, "unit": "updates/s", "ms_per_step": ms, "steps": steps,
                  "samples_per_s": batch * 1e3 / ms, "step_tflops": flops / (ms * 1e-3) / 1e12, "kernels_us": kernels,
                  "event_floor_us": floor_us}


# ----------------------------------------------------------------------------------------------- other configs
def _event_time(torch, fn, n, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(warm + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n          # ms per call


def run_config1(args, torch, dev):
    """Config 1: CartPole-v0, SimpleNetwork 4-20-20-2, CONTROL preset (scripts/dqn_params.sh), minibatch 32 and 100.
    `value`: graph-replayed updates/s (gather from the ring + fused MLP forward/backward + Adam) at B=32;
    `e2e`: the real agent loop on the cart-pole environment (env.step on the host, one update per env step)."""
    import _cabi as c
    import envs
    import network
    import replay_memory
    from dqn_agent import DQNAgent
    from network import Session

    def cfg_for(batch):
        return argparse.Namespace(
            network="simple", optimizer="adam", lr=0.001, reg_param=0.001, momentum=0.9, rmsprop_decay=0.95,
            minibatch_size=batch, max_batch=batch, num_episodes=10000, max_steps_per_episode=200,
            replay_memory_capacity=50000, target_update_freq=3000, reward_discount=0.9, init_random_action_prob=0.5,
            min_random_action_prob=0.1, random_action_explore_steps=50000, frames_per_state=1, resize_width=0,
            resize_height=0, update_freq=1, replay_start_size=0, summary_freq=100, loss="mse", seed=0, input_dtype=None,
            device_index=dev.index, cuda_graph=True, precision="fp32", logdir=tempfile.mkdtemp(prefix="dqn_bench_"))
    env = envs.make("CartPole-v0", seed=0)
    rm = replay_memory.ReplayMemory(50000, device=dev.index)
    rng = np.random.default_rng(0)
    s = env.reset()
    for _ in range(20000):                       # real cart-pole transitions under the uniform-random policy
        a = int(rng.integers(0, 2))
        s2, r, done, _ = env.step(a)
        rm.add(s, a, r, s2, done)
        s = env.reset() if done else s2
    steps = args.steps or 5000
    res = {}
    side = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(side):
        for batch in (32, 100):
            net = network.Network.create_network(cfg_for(batch), (4,), 2)
            random.seed(3)
            tab = torch.tensor([random.sample(range(rm.size()), batch) for _ in range(512)], dtype=torch.int32, device=dev)
            d_idx = torch.empty(batch, dtype=torch.int32, device=dev)
            d_idx.copy_(tab[0])
            for _ in range(3):
                net.train_step(rm, batch, d_indices=d_idx)
            n0 = c.lib.dqn_launch_count()
            net.train_step(rm, batch, d_indices=d_idx)
            launches = c.lib.dqn_launch_count() - n0
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                net.train_step(rm, batch, d_indices=d_idx)

            def one(i):
                d_idx.copy_(tab[i % 512], non_blocking=True)
                g.replay()
            ms = _event_time(torch, one, steps, warm=20)
            res[batch] = {"value": 1e3 / ms, "unit": "updates/s", "us_per_step": ms * 1e3, "launches_per_step": int(launches)}
            del g, net
    # end to end: the agent loop itself (epsilon-greedy act on the device, env.step on the host, add, update)
    cfg = cfg_for(32)
    random.seed(0)
    net = network.Network.create_network(cfg, (4,), 2)
    agent = DQNAgent(envs.make("CartPole-v0", seed=1), net, Session(), replay_memory.ReplayMemory(50000, device=dev.index),
                     cfg, enable_summary=True)
    agent.train(5, 200)
    torch.cuda.synchronize()
    u0, t0 = agent.training_steps, time.perf_counter()
    agent.train(60, 200)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    e2e_updates = agent.training_steps - u0
    # CPU baseline: the oracle's transcription of the reference step on the same network
    from oracle.replay_ref import ReplayRef
    from oracle.torch_ref import CpuBaselineAgent
    import torch as _t
    cores = len(os.sched_getaffinity(0))
    _t.set_num_threads(cores)
    ref_rm = ReplayRef(50000)
    env2 = envs.make("CartPole-v0", seed=0)
    s = env2.reset()
    for _ in range(20000):
        a = int(rng.integers(0, 2))
        s2, r, done, _ = env2.step(a)
        ref_rm.add(s, a, r, s2, done)
        s = env2.reset() if done else s2
    ref = CpuBaselineAgent("simple", (4,), 2, "adam", 0.001, 0.001, 0.9, target_update_freq=3000, seed=0)
    for _ in range(20):
        ref.train_minibatch(ref_rm, 32)
    t1 = time.perf_counter()
    ncpu = 2000
    for _ in range(ncpu):
        ref.train_minibatch(ref_rm, 32)
    cpu_v = ncpu / (time.perf_counter() - t1)
    return {
        "metric": "CartPole MLP minibatch updates/sec", "value": res[32]["value"], "unit": "updates/s", "n_gpus": 1,
        "steps": steps, "warmup": 20, "ms_per_step": res[32]["us_per_step"] / 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "real CartPole-v0 transitions (built-in cart-pole physics)",
        "config": {"workload": "CartPole-v0, SimpleNetwork 4-20-20-2, Adam lr 1e-3, gamma 0.9, 50k-transition ring, minibatch 32 "
                               "(and 100 under minibatch_100)", "minibatch_per_gpu": 32,
                   "cache": "launch/latency-bound: three small kernels per update, no roofline fraction claimed"},
        "minibatch_100": res[100], "launches_per_step": res[32]["launches_per_step"],
        "gpu_launches": res[32]["launches_per_step"] * steps,
        "e2e": {"value": e2e_updates / dt, "unit": "updates/s", "h2d_bytes_per_step": 32 + 32 + 32 * 4,
                "d2h_bytes_per_step": 4, "steps": e2e_updates,
                "path": "DQNAgent.train on the cart-pole env: _pick_action (device forward when greedy), env.step on the host, "
                        "ReplayMemory.add, _train_minibatch every env step"},
        "roofline": None,
        "cpu_baseline": {"value": cpu_v, "unit": "updates/s", "cores": cores, "kind": "port",
                         "sample": "%d updates of batch 32 through oracle/torch_ref.py + oracle/replay_ref.py" % ncpu},
    }


def run_config4(args, torch, dev):
    """Config 4: ReplayMemory sample + gather sweep on a 1M-transition uint8 ring, k = 32 ... 8192."""
    import _cabi as c
    import replay_memory
    rm = _make_ring(args, torch, c, replay_memory, 0, dev.index, dev)
    ks = (32, 64, 128, 256, 512, 1024, 2048, 4096, 8192)
    n0 = c.lib.dqn_launch_count()
    g = gather_sweep(rm, dev, ks=ks, iters=100)
    launches = c.lib.dqn_launch_count() - n0
    hbm, _, _, src = peaks()
    # end to end through the reference-facing call: ReplayMemory.get_minibatch(k) -- host index draw, H2D indices, gather,
    # both stacks + scalars copied back to host memory as the reference's tuples hold them
    k = 512
    random.seed(5)
    for _ in range(2):
        rm.get_minibatch(k)
    t0 = time.perf_counter()
    n = 20
    for _ in range(n):
        nt, t = rm.get_minibatch(k)
        cnt = sum(1 for _ in nt) + sum(1 for _ in t)
    dt = (time.perf_counter() - t0) / n
    assert cnt == k
    # CPU baseline: the reference's own ReplayMemory (restated) + the float32 feed conversion of both stacks
    from oracle.replay_ref import ReplayRef
    rng = np.random.default_rng(0)
    pool = [rng.integers(0, 256, (84, 84, 4), dtype=np.uint8) for _ in range(4096)]
    ref = ReplayRef(1000000)
    for i in range(1000000):
        ref.add(pool[i % 4096], 0, 0.0, pool[(i + 1) % 4096], False)
    t1 = time.perf_counter()
    ncpu = 5
    for _ in range(ncpu):
        nt, t = ref.get_minibatch(k)
        items = list(nt) + list(t)
        np.asarray([x[0] for x in items], dtype=np.float32)
        np.asarray([x[3] for x in items], dtype=np.float32)
    cpu_dt = (time.perf_counter() - t1) / ncpu
    top = g
    return {
        "metric": "replay gather GB/s", "value": top["achieved_gbs"], "unit": "GB/s", "n_gpus": 1, "steps": 100, "warmup": 3,
        "ms_per_step": top["us_per_launch"] / 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": "ReplayMemory sample+gather: 1M-transition uint8 84x84 frame ring, both 84x84x4 stacks + "
                               "action/reward/terminal per sampled transition, k=8192 (sweep k=32..8192 under gather.sweep)",
                   "k": 8192, "cache": "inputs larger than L2: random transitions from a 14.1 GB frame ring"},
        "gather": g, "gpu_launches": int(launches),
        "roofline": {"kernel": "gather_stack4_u8_kernel", "bound": "hbm", "achieved": top["achieved_gbs"], "peak": hbm,
                     "unit": "GB/s", "frac": top["achieved_gbs"] / hbm, "traffic": None, "peak_source": src,
                     "algorithmic_bytes_per_transition": GATHER_BYTES},
        "e2e": {"value": GATHER_BYTES * k / dt / 1e9, "unit": "GB/s", "k": k, "h2d_bytes_per_step": 4 * k,
                "d2h_bytes_per_step": k * (2 * 4 * FRAME + 13), "ms_per_call": dt * 1e3,
                "path": "ReplayMemory.get_minibatch(512): random.sample on the host, H2D indices, gather kernel, D2H of both "
                        "stacks and the scalars, python tuples (the reference's return type)"},
        "cpu_baseline": {"value": GATHER_BYTES * k / cpu_dt / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                         "sample": "%d x (get_minibatch(512) on a 1M-entry deque + float32 conversion of both stack lists), "
                                   "oracle/replay_ref.py" % ncpu, "ms_per_call": cpu_dt * 1e3},
    }


def run_config5(args, torch, dist, dp, rank, local, world, dev):
    """Config 5: Breakout shape (4 actions): batched greedy act for 1024 synthetic env states (split over the GPUs),
    one B=512/GPU train step, one target sync -- each timed, and their sum."""
    import _cabi as c
    import network
    import replay_memory
    A, batch, envs_total = 4, 512, 1024
    per_gpu = envs_total // world
    cfg = atari_config(batch, args.arch, args.precision)
    cfg.device_index = local
    net = network.Network.create_network(cfg, (84, 84, 4), A, num_replicas=world)
    dp.setup_network(net, cfg, mode=args.dp)
    side = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(side):
        rm = _make_ring(args, torch, c, replay_memory, rank, local, dev)
        gen = torch.Generator(device=dev).manual_seed(2 + rank)
        states = torch.randint(0, 256, (per_gpu, 84, 84, 4), dtype=torch.uint8, device=dev, generator=gen)
        random.seed(9 + rank)
        tab = torch.tensor([random.sample(range(rm.size()), batch) for _ in range(64)], dtype=torch.int32, device=dev)
        d_idx = torch.empty(batch, dtype=torch.int32, device=dev)
        d_idx.copy_(tab[0])
        dp.train_step(net, rm, batch, d_idx)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            dp.train_step(net, rm, batch, d_idx)
        n = args.steps or 100

        def act(i):
            net.predict(states, argmax=True)

        def train(i):
            d_idx.copy_(tab[i % 64], non_blocking=True)
            g.replay()

        def sync(i):
            net.sync_target()

        def cycle(i):
            act(i); train(i); sync(i)
        out = {}
        for name, fn in (("act", act), ("train", train), ("target_sync", sync), ("cycle", cycle)):
            if world > 1:
                torch.cuda.synchronize(); dist.barrier()
            ms = _event_time(torch, fn, n, warm=5)
            t = torch.tensor([ms], device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out[name] = float(t.item()) * 1e3
        del g
    if rank != 0:
        return None
    return {
        "metric": "act(1024 envs) + train step (B=512/GPU) + target sync cycles/sec", "value": 1e6 / out["cycle"],
        "unit": "cycles/s", "n_gpus": world, "steps": n, "warmup": 5, "ms_per_step": out["cycle"] / 1e3,
        "higher_is_better": True, "scaling": "strong for the act phase (1024 envs split over the GPUs), weak for the update",
        "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
        "config": {"workload": "Breakout shape (4 actions): batched greedy act for 1024 synthetic 84x84x4 uint8 env states "
                               "(%d per GPU) + one minibatch-512/GPU update + target-network sync" % per_gpu,
                   "parallelism": "dp%d" % world, "cache": "update gathers fresh random transitions from a 14.1 GB ring"},
        "phases_us": out, "sum_of_phases_us": out["act"] + out["train"] + out["target_sync"],
        "roofline": None, "cpu_baseline": None, "e2e": None, "dp_exchange": (args.dp if world > 1 else None),
    }


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


def _traffic(name, batch):
    """DRAM bytes per launch of `name` from the committed ncu --set full captures, else None.  Not measured in this run
    (that needs a profiler): the source file is named beside the value."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        v = t["dram_bytes_per_launch"].get(str(batch), {}).get(name)
        return v, (t.get("source") if v is not None else None)
    except Exception:
        return None, None


def roofline(kernels, arch, batch, nparams, floor_us=0.0):
    """Roofline of the dominant kernel of the step (largest average duration on the critical path).  Durations come
    from CUDA-event brackets (see _kernel_times); `floor_us` (the same bracket around an empty launch) is subtracted for
    eager brackets and 0 for in-graph ones."""
    if not kernels:
        return None
    hbm, burst, sustained, src = peaks()
    # only kernels on the critical path of the step (the head / fc / conv3 / conv2 weight gradients and the loss run on
    # forked streams in the shadow of the data-gradient chain)
    side = ("head_wgrad", "loss", "wgrad_fc", "wgrad_conv3", "wgrad_conv2")
    crit = {k: v for k, v in kernels.items() if k not in side} or kernels
    name = max(crit, key=lambda k: crit[k])
    raw_us = crit[name]
    us = max(raw_us - floor_us, 0.25 * raw_us)
    macs = MACS[arch]
    layer = {"conv1": 0, "conv2": 1, "conv3": 2, "fc": 3}
    traffic, tsrc = _traffic(name, batch)
    if name.startswith(("fwd_", "wgrad_", "dgrad_")):
        l = layer[name.split("_", 1)[1]]
        passes = 2 if name.startswith("fwd_") else 1          # online + target forward share one launch
        flops = 2.0 * macs[l] * batch * passes
        ach = flops / (us * 1e-6) / 1e12
        return {"kernel": name, "bound": "tensor", "achieved": ach, "peak": sustained, "unit": "TFLOP/s",
                "frac": ach / sustained, "traffic": traffic, "traffic_source": tsrc, "us_per_launch": us,
                "us_event_bracket": raw_us, "event_floor_us": floor_us, "minibatch": batch,
                "peak_source": src + " bf16 sustained",
                "note": "implicit GEMM against the measured bf16 tensor-pipe peak; at batch 32 every kernel is a few "
                        "CTAs of latency-bound work and the fraction only says so -- see roofline_large_batch"}
    if name == "fc_td_fused":
        # fc forward of both nets + Q head of both + fc data gradient, one cluster launch
        flops = 2.0 * (3 * macs[3] + 2 * 256 * A_PONG) * batch
        ach = flops / (us * 1e-6) / 1e12
        return {"kernel": name, "bound": "tensor", "achieved": ach, "peak": sustained, "unit": "TFLOP/s",
                "frac": ach / sustained, "traffic": traffic, "traffic_source": tsrc, "us_per_launch": us,
                "us_event_bracket": raw_us, "event_floor_us": floor_us, "minibatch": batch,
                "peak_source": src + " bf16 sustained",
                "note": "fc forward (both nets) + Q head + TD target/loss/gradient + fc data gradient in ONE 8-CTA cluster "
                        "launch: 13 MFLOP behind three cluster barriers, a latency chain, not a throughput kernel -- the "
                        "fraction only says so; see roofline_large_batch for a kernel that fills the GPU"}
    if name == "gather":
        byts = GATHER_BYTES * batch
    elif name in ("optimizer", "reduce_grad"):
        byts = 20.0 * nparams
    else:
        byts = 0.0
    ach = byts / (us * 1e-6) / 1e9
    return {"kernel": name, "bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm,
            "traffic": traffic, "traffic_source": tsrc, "us_per_launch": us, "us_event_bracket": raw_us,
            "event_floor_us": floor_us, "minibatch": batch, "peak_source": src}


def gather_sweep(rm, dev, ks=(8192,), iters=30):
    """Replay gather GB/s (config 4 of BASELINE.json) per k, CUDA-event timed, fresh random indices each launch (the
    frame ring is far larger than L2).  Returns the k = max(ks) entry with the others under "sweep"."""
    import torch
    import _cabi as c
    size = rm.size()
    hbm, _, _, src = peaks()
    st = c.current_stream()
    out = []
    for k in ks:
        random.seed(77 + k)
        n = iters if k >= 512 else 200
        tables = torch.tensor([random.sample(range(size), k) for _ in range(n + 3)], dtype=torch.int32, device=dev)
        states = torch.empty((k, 84, 84, 4), dtype=torch.uint8, device=dev)
        nexts = torch.empty_like(states)
        acts = torch.empty(k, dtype=torch.int32, device=dev)
        rews = torch.empty(k, dtype=torch.float64, device=dev)
        terms = torch.empty(k, dtype=torch.uint8, device=dev)

        def go(i):
            c.check(c.lib.dqn_replay_gather(rm.handle, tables[i].data_ptr(), k, states.data_ptr(), nexts.data_ptr(),
                                            acts.data_ptr(), rews.data_ptr(), terms.data_ptr(), st))
        for i in range(3):
            go(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(3, 3 + n):
            go(i)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / n * 1e3
        gbs = GATHER_BYTES * k / (us * 1e-6) / 1e9
        out.append({"k": k, "us_per_launch": us, "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / hbm})
        del tables, states, nexts
    top = dict(max(out, key=lambda e: e["k"]))
    top.update({"peak_gbs": hbm, "bytes_per_transition": GATHER_BYTES, "peak_source": src, "sweep": out,
                "timing": "back-to-back launches between two CUDA events (launch overhead overlaps the previous kernel)"})
    return top


def run_e2e(args, cfg, net, rm, world, dev, DQNAgent, FrameBuffer, dist):
    """updates/s through the public API with HOST inputs, the same loop the reference arm times on the CPU: per step
    update_freq x ReplayMemory.add(state, action, reward, next_state, terminal) with host uint8 stacks (the reference's
    five-argument signature) and one DQNAgent._train_minibatch (host random.sample -> H2D indices -> CUDA graph -> loss
    read back to the host).  Also reported: the same with the actor-side FrameBuffer.append/get_state in the loop (the
    agent's own episode loop, which hands add() the newest frame as a contiguous array)."""
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
    NP = 256
    pool = rng.integers(0, 256, (NP, 84, 84), dtype=np.uint8)
    # host stacks as an actor would hold them: stack i = frames i-3..i (HWC, channel = time), so that stack i+1 is stack i
    # shifted by one frame -- what ReplayMemory.add receives from the reference's episode loop (dqn_agent.py:96-100)
    stacks = [np.ascontiguousarray(np.stack([pool[(i - 3 + c) % NP] for c in range(4)], axis=-1)) for i in range(NP)]
    n = args.e2e_steps
    warm = 10
    fi = 0

    def one():
        nonlocal fi
        for _ in range(cfg.update_freq):
            rm.add(stacks[fi % NP], fi % A_PONG, float(fi % 3 - 1), stacks[(fi + 1) % NP], False)
            fi += 1
        agent._train_minibatch(batch)          # summary_freq=1 -> the loss of every update is read back to the host
        return agent.last_loss                 # (asynchronously: it is picked up while the next update is prepared)

    fb = FrameBuffer(4)
    fb.append(pool[0])
    fstate = [fb.get_state()]

    def one_fb():
        nonlocal fi
        state = fstate[0]
        for _ in range(cfg.update_freq):
            fb.append(pool[fi % NP]); fi += 1
            nxt = fb.get_state()
            rm.add(state, fi % A_PONG, float(fi % 3 - 1), nxt, False, newest=fb.frames[-1])     # as DQNAgent._record does
            state = nxt
        fstate[0] = state
        agent._train_minibatch(batch)
        return agent.last_loss

    def timed(fn, count):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(count):
            fn()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([dt], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return dt

    dt = timed(one, n)
    dt_fb = timed(one_fb, max(n // 4, 50))
    agent.flush_summaries()
    loss = agent.last_loss
    assert loss is not None and np.isfinite(loss)
    agent._graphs.clear()
    # the step graph copies a fixed prefix of the pinned staging buffer: 8 frame slots + header + 8 records, and the indices
    gi = rm.GRAPH_INGEST_FRAMES
    h2d = gi * FRAME + 64 + gi * 32 + batch * 4
    return {"value": world * n / dt, "unit": "updates/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
            "steps": n, "ms_per_step": dt / n * 1e3,
            "with_framebuffer": {"value": world * max(n // 4, 50) / dt_fb, "unit": "updates/s",
                                 "path": "FrameBuffer.append/get_state + ReplayMemory.add(..., newest=frame) x%d + "
                                         "_train_minibatch" % cfg.update_freq},
            "path": "ReplayMemory.add x%d (host uint8 stacks, reference signature) + DQNAgent._train_minibatch "
                    "(random.sample on host; one CUDA-graph launch per update moves the staged frames into the ring, "
                    "updates straight from the ring and copies the loss to pinned host memory, where it is collected one "
                    "update later)" % cfg.update_freq}


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
