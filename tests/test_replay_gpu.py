"""GPU parity of the replay ring + gather (through the C ABI) against fixtures recorded from the
reference's own ReplayMemory/FrameBuffer, and against the oracle on random scripts."""
import os
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _mods():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import frame_buffer
    import replay_memory
    return replay_memory, frame_buffer


@pytest.mark.parametrize("name", ["small", "atari", "vec"])
def test_episode_script_bit_exact_vs_reference_fixture(golden_dir, name):
    rm_mod, fb_mod = _mods()
    from oracle.gen_golden import episode_script
    g = np.load(os.path.join(golden_dir, "episode_script_%s.npz" % name))
    batches, size = episode_script(rm_mod.ReplayMemory, fb_mod.FrameBuffer, g["frames"], list(g["ep_lens"]),
                                   list(g["ep_done"]), int(g["capacity"][0]), int(g["k"][0]),
                                   int(g["sample_every"][0]), int(g["seed"][0]), int(g["frames_per_state"][0]))
    assert size == int(g["final_size"][0]) and len(batches) == int(g["nbatches"][0])
    for i, b in enumerate(batches):
        for key, v in b.items():
            ref = g["b%d_%s" % (i, key)]
            got = np.asarray(v)
            assert got.shape == ref.shape and got.dtype == ref.dtype, (i, key, got.dtype, ref.dtype)
            assert np.array_equal(got, ref), (i, key)


def test_sample_index_stream_matches_reference_fixture(golden_dir):
    rm_mod, _ = _mods()
    g = np.load(os.path.join(golden_dir, "replay_sample_indices.npz"))
    for ci in range(int(g["ncases"][0])):
        adds, cap, k, seed = (int(v) for v in g["case%d_cfg" % ci])
        if adds > 20000:
            continue                      # the big cases are covered on the CPU by the oracle test
        rm = rm_mod.ReplayMemory(cap)
        for i in range(adds):
            rm.add(np.array([i], np.int64), i % 6, float(i % 3 - 1), np.array([i + 1], np.int64), (i % 11) == 10)
        assert [rm.size(), rm.capacity()] == list(g["case%d_size" % ci])
        random.seed(seed)
        nt, t = rm.get_minibatch(k)
        if "case%d_none" % ci in g.files:
            assert nt is None and t is None
            continue
        nt, t = list(nt), list(t)
        assert [int(x[0][0]) for x in nt] == list(g["case%d_nt" % ci])
        assert [int(x[0][0]) for x in t] == list(g["case%d_t" % ci])
        assert [int(v[0]) for v in rm_mod.ReplayMemory.get_next_states(nt)] == list(g["case%d_next_nt" % ci])
        assert [int(v[0]) for v in rm_mod.ReplayMemory.get_states(t)] == list(g["case%d_state_t" % ci])
        assert all(x[4] is False for x in nt) and all(x[4] is True for x in t)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_random_scripts_vs_oracle_including_unchained_stacks(seed):
    """Arbitrary callers: stacks that do not chain, wrap-around, sampling interleaved with adds."""
    rm_mod, _ = _mods()
    from oracle.replay_ref import ReplayRef
    rng = np.random.default_rng(seed)
    cap = int(rng.integers(5, 40))
    S = int(rng.choice([2, 3, 4, 5]))
    shape = (int(rng.integers(2, 9)), int(rng.integers(2, 9)) * (4 if seed == 0 else 1))
    mine, ref = rm_mod.ReplayMemory(cap), ReplayRef(cap)
    prev = None
    for step in range(4 * cap + 7):
        mode = rng.integers(0, 10)
        if prev is None or mode == 0:                       # unrelated state
            state = rng.integers(0, 256, shape + (S,), dtype=np.uint8)
        else:
            state = prev
        if mode == 1:                                       # next state not a shift of state
            nxt = rng.integers(0, 256, shape + (S,), dtype=np.uint8)
        else:
            nxt = np.concatenate([state[..., 1:], rng.integers(0, 256, shape + (1,), dtype=np.uint8)], axis=-1)
        a, r, t = int(rng.integers(0, 6)), float(rng.standard_normal()), bool(rng.random() < 0.2)
        mine.add(state, a, r, nxt, t)
        ref.add(state, a, r, nxt, t)
        prev = nxt
        assert mine.size() == ref.size() == mine.device_size()
        if step % 3 == 0:
            k = int(rng.integers(1, 8))
            st = random.getstate()
            random.seed(1000 + step)
            got = mine.get_minibatch(k)
            random.seed(1000 + step)
            exp = ref.get_minibatch(k)
            random.setstate(st)
            if exp[0] is None:
                assert got == (None, None)
                continue
            for gi, ei in zip(got, exp):
                gi, ei = list(gi), list(ei)
                assert len(gi) == len(ei)
                for x, y in zip(gi, ei):
                    assert np.array_equal(x[0], y[0]) and x[1] == y[1] and x[2] == y[2]
                    assert np.array_equal(x[3], y[3]) and x[4] == y[4]


def test_large_gather_property_vs_torch_indexing():
    """Full-size shapes: 84x84 frames appended from device memory, k=4096 stacks; every channel of
    every gathered stack must equal the ring frame it came from (checked with torch indexing)."""
    rm_mod, _ = _mods()
    import torch
    import _cabi as c
    n_frames, cap, k = 20000, 12000, 4096
    dev = torch.device("cuda", torch.cuda.current_device())
    gen = torch.Generator(device=dev).manual_seed(0)
    frames = torch.randint(0, 256, (n_frames, 84, 84), dtype=torch.uint8, device=dev, generator=gen)
    h = c.c_vp()
    c.check(c.lib.dqn_replay_create(c.C.byref(h), cap, 0, 84 * 84, 1, 4, dev.index))
    pos = c.c_i64()
    c.check(c.lib.dqn_replay_append_frames_device(h, frames.data_ptr(), n_frames, c.C.byref(pos), c.current_stream()))
    assert pos.value == 0
    for t in range(cap + 500):                      # wraps the record ring by 500
        p = 3 + t
        c.check(c.lib.dqn_replay_add(h, p, p + 1, t % 6, float(t % 5) - 2.0, 1 if t % 97 == 0 else 0))
    size = c.c_i64()
    c.lib.dqn_replay_size(h, c.C.byref(size), None)
    assert size.value == cap
    random.seed(5)
    idx = random.sample(range(cap), k)
    d_idx = torch.tensor(idx, dtype=torch.int32, device=dev)
    states = torch.empty((k, 84, 84, 4), dtype=torch.uint8, device=dev)
    nexts = torch.empty_like(states)
    acts = torch.empty(k, dtype=torch.int32, device=dev)
    rews = torch.empty(k, dtype=torch.float64, device=dev)
    terms = torch.empty(k, dtype=torch.uint8, device=dev)
    c.check(c.lib.dqn_replay_gather(h, d_idx.data_ptr(), k, states.data_ptr(), nexts.data_ptr(), acts.data_ptr(),
                                    rews.data_ptr(), terms.data_ptr(), c.current_stream()))
    torch.cuda.synchronize()
    t_of = torch.tensor(idx, device=dev) + 500                    # logical 0 = transition 500
    p = 3 + t_of
    for ch in range(4):
        assert torch.equal(states[..., ch], frames[p - 3 + ch])
        assert torch.equal(nexts[..., ch], frames[p - 2 + ch])
    assert torch.equal(acts.long(), t_of % 6)
    assert torch.equal(rews, (t_of % 5).double() - 2.0)
    assert torch.equal(terms.long(), (t_of % 97 == 0).long())
    # gathering only one of the outputs is allowed
    c.check(c.lib.dqn_replay_gather(h, d_idx.data_ptr(), k, None, nexts.data_ptr(), None, None, None, c.current_stream()))
    torch.cuda.synchronize()
    assert torch.equal(nexts[..., 3], frames[p + 1])
    c.lib.dqn_replay_destroy(h)


def test_frame_ring_overflow_and_bad_arguments_fail_loudly():
    rm_mod, _ = _mods()
    import _cabi as c
    h = c.c_vp()
    assert c.lib.dqn_replay_create(c.C.byref(h), 4, 0, 16, 1, 0, 0) != 0
    assert "Frames per state" in c.last_error()
    c.check(c.lib.dqn_replay_create(c.C.byref(h), 4, 9, 16, 1, 2, 0))
    fr = np.zeros((9, 16), np.uint8)
    pos = c.c_i64()
    c.check(c.lib.dqn_replay_append_frames(h, fr.ctypes.data, 9, c.C.byref(pos)))
    c.check(c.lib.dqn_replay_add(h, 1, 2, 0, 0.0, 0))
    assert c.lib.dqn_replay_append_frames(h, fr.ctypes.data, 2, c.C.byref(pos)) == -4      # would overwrite frame 0
    assert "overflow" in c.last_error()
    assert c.lib.dqn_replay_add(h, 0, 1, 0, 0.0, 0) == -1                                 # stack reaches below frame 0
    c.lib.dqn_replay_destroy(h)
    rm = rm_mod.ReplayMemory(8)
    with pytest.raises(RuntimeError):
        rm.sample(3)


def test_lock_step_actor_batch_add_equals_per_env_framebuffers_and_adds():
    """E interleaved environments through ReplayMemory.add_batch (one device frame per environment per
    call) against E FrameBuffers + one reference-semantics replay memory fed in environment order -- bit-exact, through
    episode starts (dup-filled windows), terminals, eviction and wrap-around of the environments' frame lanes."""
    _mods()
    import torch
    from frame_buffer import FrameBuffer
    from oracle.replay_ref import ReplayRef
    from replay_memory import ReplayMemory
    E, S, cap, steps = 5, 4, 64, 60
    rng = np.random.default_rng(11)
    rm = ReplayMemory(cap, frame_capacity=E * 64)          # 64 slots per lane: every lane wraps during the run (~110 frames each)
    rm.set_lanes(E)
    ref = ReplayRef(cap)
    fbs = [FrameBuffer(S) for _ in range(E)]
    states = [None] * E
    left = [0] * E                                          # steps until the environment's episode ends
    for step in range(steps):
        frames = rng.integers(0, 256, (E, 84, 84), dtype=np.uint8)
        a = rng.integers(0, 6, E).astype(np.int32)
        r = rng.integers(-1, 2, E).astype(np.float64)
        t = np.zeros(E, np.uint8)
        s = np.zeros(E, np.uint8)
        for e in range(E):
            if left[e] == 0:                                # env.reset(): FrameBuffer.clear + append, no transition
                fbs[e].clear()
                fbs[e].append(frames[e])
                states[e] = fbs[e].get_state()
                left[e] = int(rng.integers(1, 9))
                s[e] = 1
            else:
                left[e] -= 1
                t[e] = 1 if left[e] == 0 else 0
                fbs[e].append(frames[e])
                nxt = fbs[e].get_state()
                ref.add(states[e], int(a[e]), float(r[e]), nxt, bool(t[e]))
                states[e] = nxt
        rm.add_batch(torch.from_numpy(frames).cuda(), a, r, t, s)
        cur = rm.current_states().cpu().numpy()
        for e in range(E):
            assert np.array_equal(cur[e], states[e]), (step, e)
    assert rm.size() == ref.size() == cap                  # the oldest transitions were evicted like deque(maxlen)
    b = rm.sample(cap, indices=list(range(cap)))
    st, nx = b.states.cpu().numpy(), b.next_states.cpu().numpy()
    ac, rw, tm = b.actions.cpu().numpy(), b.rewards.cpu().numpy(), b.terminals.cpu().numpy()
    for i, (s0, a0, r0, s1, t0) in enumerate(ref.memory):
        assert np.array_equal(st[i], s0) and np.array_equal(nx[i], s1), i
        assert ac[i] == a0 and rw[i] == r0 and bool(tm[i]) == t0, i


def test_batched_actor_epsilon_greedy_matches_the_reference_rule():
    """BatchedActor: epsilon decays before every roll, one roll per environment in environment order on Python's global
    generator, greedy actions from ONE batched forward; an agent-style loop over 6 environments stays consistent with the
    ring (sizes, finite Q-values)."""
    _mods()
    import envs
    import network
    from preprocess import BatchedActor
    from replay_memory import ReplayMemory
    from gpu_helpers import make_config
    E = 6
    cfg = make_config(network='cnn', minibatch_size=8, max_batch=8, init_random_action_prob=0.9, min_random_action_prob=0.1,
                      random_action_explore_steps=50)
    net = network.Network.create_network(cfg, (84, 84, 4), 6)
    rm = ReplayMemory(500)
    rm.set_lanes(E)
    es = [envs.SyntheticAtariEnv('SyntheticPong-v0', num_actions=6, seed=e, frame_shape=(84, 84), min_len=5, max_len=12) for e in range(E)]
    actor = BatchedActor(es, net, rm, cfg)
    random.seed(4)
    mirror = random.Random(4)
    eps = cfg.init_random_action_prob
    for _ in range(30):
        nstep = sum(1 for x in actor._needs_reset if not x)
        actor.step()
        for e in range(nstep):                               # one decay, then one roll, per stepping environment
            if eps > cfg.min_random_action_prob:
                eps -= (cfg.init_random_action_prob - cfg.min_random_action_prob) / cfg.random_action_explore_steps
            mirror.random()
        assert abs(actor.random_action_prob - eps) < 1e-12
        assert random.getstate() == mirror.getstate()
    assert 0 < rm.size() <= 30 * E and actor.random_action_prob <= cfg.min_random_action_prob + 1e-9
    q = net.predict(rm.current_states())
    assert q.shape == (E, 6) and bool(q.isfinite().all())


def test_index_table_next_walks_the_rows_and_wraps():
    """dqn_index_table_next: the captured update's own source of sample indices -- row (cursor % rows) of a
    device-resident table, cursor advanced on the device, also inside a CUDA graph."""
    _mods()
    import torch
    import replay_memory as R
    rows = [[r * 10 + j for j in range(5)] for r in range(3)]
    tab = R.IndexTable(rows, torch.device("cuda", 0))
    d_idx = torch.full((5,), -1, dtype=torch.int32, device="cuda")
    seen = []
    for _ in range(4):
        tab.next(d_idx)
        seen.append(d_idx.cpu().tolist())
    assert seen == [rows[0], rows[1], rows[2], rows[0]] and int(tab.cursor.item()) == 4
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            tab.next(d_idx)
        g.replay(); g.replay()
    torch.cuda.synchronize()
    assert d_idx.cpu().tolist() == rows[(4 + 2 - 1) % 3] and int(tab.cursor.item()) == 6
