"""Generates tests/golden/*.npz (TEST INFRASTRUCTURE; run in the build container only).

    python -m oracle.gen_golden            # needs /root/reference (read-only); writes tests/golden/

Data half: imports the reference's OWN modules from /root/reference/src (replay_memory.py,
frame_buffer.py, utils.py import fine without TensorFlow) and records what they return on
seeded scripts -> these fixtures pin oracle/replay_ref.py and the CUDA replay ring.
Model half: TensorFlow cannot run here, so the net fixtures come from oracle/net_ref.py
itself (float64) and only guard it against regressions -- "parity unpinned".
"""
from __future__ import annotations

import os
import random
import sys

import numpy as np

REF_SRC = "/root/reference/src"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _ref():
    sys.path.insert(0, REF_SRC)
    import replay_memory as ref_rm           # noqa: E402  (the reference's, unmodified)
    import frame_buffer as ref_fb            # noqa: E402
    import utils as ref_utils                # noqa: E402
    sys.path.pop(0)
    assert ref_rm.__file__.startswith(REF_SRC)
    return ref_rm, ref_fb, ref_utils


# ------------------------------------------------------------------ 1. index stream of get_minibatch
def gen_sample_indices(ref_rm):
    """Payloads are ints, so the record is exactly which logical transitions random.sample picks."""
    cases = [  # (adds, capacity, k, seed)
        (100, 100, 32, 0), (277, 1000, 32, 1), (278, 1000, 32, 2), (50000, 50000, 100, 3),
        (1200, 500, 32, 4),            # wrapped ring: live items are adds 700..1199
        (8, 5, 3, 5), (64, 64, 64, 6), (20, 50, 32, 7),   # last: size < k -> (None, None)
        (100000, 100000, 512, 8),
    ]
    out = {}
    for ci, (adds, cap, k, seed) in enumerate(cases):
        rm = ref_rm.ReplayMemory(cap)
        for i in range(adds):
            rm.add(i, i % 6, float(i % 3 - 1), i + 1, (i % 11) == 10)
        random.seed(seed)
        nt, t = rm.get_minibatch(k)
        out["case%d_cfg" % ci] = np.array([adds, cap, k, seed], np.int64)
        out["case%d_size" % ci] = np.array([rm.size(), rm.capacity()], np.int64)
        if nt is None:
            out["case%d_none" % ci] = np.array([1])
            continue
        nt, t = list(nt), list(t)
        out["case%d_nt" % ci] = np.array([x[0] for x in nt], np.int64)
        out["case%d_t" % ci] = np.array([x[0] for x in t], np.int64)
        out["case%d_next_nt" % ci] = np.array(list(ref_rm.ReplayMemory.get_next_states(nt)), np.int64)
        out["case%d_state_t" % ci] = np.array(list(ref_rm.ReplayMemory.get_states(t)), np.int64)
    out["ncases"] = np.array([len(cases)])
    np.savez_compressed(os.path.join(OUT, "replay_sample_indices.npz"), **out)


# ------------------------------------------------------------------ 2. FrameBuffer
def gen_framebuffer(ref_fb):
    rng = np.random.default_rng(11)
    frames = rng.integers(0, 256, (9, 6, 8), dtype=np.uint8)
    out = {"frames": frames}
    fb = ref_fb.FrameBuffer(4)
    assert fb.get_state() is None
    for i in range(6):
        fb.append(frames[i])
        out["fb4_after%d" % i] = fb.get_state()
    fb.clear()
    out["fb4_cleared_is_none"] = np.array([fb.get_state() is None])
    fb.append(frames[6])
    out["fb4_restart"] = fb.get_state()
    st = fb.get_state()
    out["fb4_meta"] = np.array([st.shape[0], st.shape[1], st.shape[2], st.strides[0], st.strides[1], st.strides[2],
                                int(st.flags["C_CONTIGUOUS"])], np.int64)
    fb1 = ref_fb.FrameBuffer(1)
    fb1.append(frames[0]); fb1.append(frames[1])
    out["fb1_state"] = fb1.get_state()                       # bare frame, not a length-1 stack
    fb2 = ref_fb.FrameBuffer(2, preprocessor=lambda f: (f // 2).astype(np.uint8))
    fb2.append(frames[2]); fb2.append(frames[3]); fb2.append(frames[4])
    out["fb2_pre_state"] = fb2.get_state()
    try:
        ref_fb.FrameBuffer(0)
        out["fb0_raises"] = np.array([0])
    except RuntimeError as e:
        out["fb0_raises"] = np.array([1])
        out["fb0_msg"] = np.array(str(e))
    np.savez_compressed(os.path.join(OUT, "framebuffer.npz"), **out)


# ------------------------------------------------------------------ 3. agent-style episode script
def episode_script(ReplayMemory, FrameBuffer, frames, ep_lens, ep_done, capacity, k, sample_every, seed,
                   frames_per_state=4):
    """Drives FrameBuffer+ReplayMemory the way DQNAgent.train_episode does (dqn_agent.py:72-106):
    clear/reset/append/get_state, then per step append -> get_state -> add(state,a,r,next_state,done);
    every ``sample_every`` adds (once size>=k) call get_minibatch(k).  Returns the sampled batches."""
    rm = ReplayMemory(capacity)
    fb = FrameBuffer(frames_per_state)
    random.seed(seed)
    fi = 0
    adds = 0
    batches = []
    for L, done_at_end in zip(ep_lens, ep_done):
        fb.clear()
        fb.append(frames[fi]); fi += 1
        state = fb.get_state()
        for step in range(L):
            action = (fi * 7 + step) % 5
            reward = float((fi % 5) - 2) * 0.5
            done = bool(done_at_end and step == L - 1)
            fb.append(frames[fi]); fi += 1
            nxt = fb.get_state()
            rm.add(state, action, reward, nxt, done)
            state = nxt
            adds += 1
            if adds % sample_every == 0 and rm.size() >= k:
                nt, t = rm.get_minibatch(k)
                nt, t = list(nt), list(t)
                items = nt + t                                   # dqn_agent.py:241 order
                batches.append(dict(
                    n_nonterminal=len(nt),
                    states=np.stack([x[0] for x in items]), actions=np.array([x[1] for x in items], np.int32),
                    rewards=np.array([x[2] for x in items], np.float64),
                    next_states=np.stack([x[3] for x in items]), terminals=np.array([x[4] for x in items], np.uint8)))
    return batches, rm.size()


def gen_episode_script(ref_rm, ref_fb):
    rng = np.random.default_rng(21)
    for name, hw, cap, k, fps in (("small", (8, 12), 40, 8, 4), ("atari", (84, 84), 24, 4, 4),
                                  ("vec", None, 30, 6, 1)):
        ep_lens = [1, 2, 3, 5, 9, 4, 17, 1, 6, 30, 2, 11, 8, 3, 20]
        ep_done = [True, True, False, True, True, False, True, True, True, False, True, True, False, True, True]
        if name == "atari":
            ep_lens, ep_done = ep_lens[:9], ep_done[:9]
        nframes = sum(ep_lens) + len(ep_lens)
        if hw is None:
            frames = rng.standard_normal((nframes, 4))             # CartPole-like float64 observations
        else:
            frames = rng.integers(0, 256, (nframes,) + hw, dtype=np.uint8)
        batches, size = episode_script(ref_rm.ReplayMemory, ref_fb.FrameBuffer, frames, ep_lens, ep_done,
                                       cap, k, 5, seed=33, frames_per_state=fps)
        out = dict(frames=frames, ep_lens=np.array(ep_lens), ep_done=np.array(ep_done), capacity=np.array([cap]),
                   k=np.array([k]), sample_every=np.array([5]), seed=np.array([33]), final_size=np.array([size]),
                   frames_per_state=np.array([fps]), nbatches=np.array([len(batches)]))
        for i, b in enumerate(batches):
            for key, v in b.items():
                out["b%d_%s" % (i, key)] = np.asarray(v)
        np.savez_compressed(os.path.join(OUT, "episode_script_%s.npz" % name), **out)


# ------------------------------------------------------------------ 4. utils
def gen_utils(ref_utils):
    out = {}
    out["one_hot_2_6"] = ref_utils.one_hot(2, 6)
    out["one_hot_0_2"] = ref_utils.one_hot(0, 2)
    out["decay_per_step"] = np.array([ref_utils.decay_per_step(1.0, 0.1, 1000000),
                                      ref_utils.decay_per_step(0.5, 0.1, 50000),
                                      ref_utils.decay_per_step(1.0, 0.1, 0),
                                      ref_utils.decay_per_step(1.0, 0.1, -3)], np.float64)
    out["decay"] = np.array([ref_utils.decay(1.0, 0.1, 0.5), ref_utils.decay(0.15, 0.1, 0.5)], np.float64)
    a, b = ref_utils.partition(lambda v: v % 3 == 0, [5, 3, 9, 1, 6, 7, 0, 2])
    out["partition_false"] = np.array(list(a)); out["partition_true"] = np.array(list(b))
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (210, 160, 3), dtype=np.uint8)
    out["resize_in"] = img
    out["resize_out_84"] = ref_utils.resize_image(img, 84, 84)
    out["resize_out_w40_h30_shape"] = np.array(ref_utils.resize_image(img, 40, 30).shape)
    np.savez_compressed(os.path.join(OUT, "utils.npz"), **out)


# ------------------------------------------------------------------ 5. net fixtures (self-generated, unpinned)
def _sample(t, n=64):
    f = np.asarray(t).ravel()
    idx = np.linspace(0, f.size - 1, min(n, f.size)).astype(np.int64)
    return np.concatenate([[f.sum(), np.abs(f).sum(), (f * f).sum()], f[idx]])


def gen_net():
    from . import net_ref as N
    out = {}
    cfgs = [("cnn", (84, 84, 4), 6, 4, "rmsprop", 0.00025, 0.99), ("cnn", (84, 84, 4), 4, 3, "adam", 0.001, 0.99),
            ("nature", (84, 84, 4), 6, 2, "rmsprop", 0.00025, 0.99), ("simple", (4,), 2, 5, "adam", 0.001, 0.9),
            ("simple", (4,), 2, 7, "sgd", 0.001, 0.9)]
    for ci, (arch, ishape, A, B, opt, lr, gamma) in enumerate(cfgs):
        rng = np.random.default_rng(100 + ci)
        # stddev 0.05 instead of the reference's 0.01 so activations/gradients are not vanishingly small
        tr = N.RefTrainer(arch, ishape, A, optimizer=opt, lr=lr, reg_param=0.001, gamma=gamma, seed=ci,
                          params=N.init_params(arch, ishape, A, seed=ci, stddev=0.05), target_update_freq=2)
        for i in range(1, len(tr.params), 2):
            tr.params[i] += 0.01 * rng.standard_normal(tr.params[i].shape)
        tr.target = [p + 0.002 * rng.standard_normal(p.shape) for p in tr.params]
        if len(ishape) == 3:
            mk = lambda: rng.integers(0, 256, (B,) + ishape, dtype=np.uint8)
        else:
            mk = lambda: rng.standard_normal((B,) + ishape)
        out["c%d_cfg" % ci] = np.array([arch, opt, str(A), str(B), repr(lr), repr(gamma), repr(ishape)])
        for step in range(3):
            s, s2 = mk(), mk()
            a = rng.integers(0, A, B)
            r = rng.integers(-1, 2, B).astype(np.float64)
            term = (rng.random(B) < 0.3)
            loss, q, grads, y = tr.train(s, a, r, s2, term)
            out["c%d_s%d_loss" % (ci, step)] = np.array([loss])
            out["c%d_s%d_q" % (ci, step)] = q
            out["c%d_s%d_y" % (ci, step)] = y
            for pi, (g, p) in enumerate(zip(grads, tr.params)):
                out["c%d_s%d_g%d" % (ci, step, pi)] = _sample(g)
                out["c%d_s%d_p%d" % (ci, step, pi)] = _sample(p)
            out["c%d_s%d_tgt0" % (ci, step)] = _sample(tr.target[0])
    np.savez_compressed(os.path.join(OUT, "net_ref_regression.npz"), **out)


def main():
    os.makedirs(OUT, exist_ok=True)
    ref_rm, ref_fb, ref_utils = _ref()
    gen_sample_indices(ref_rm)
    gen_framebuffer(ref_fb)
    gen_episode_script(ref_rm, ref_fb)
    gen_utils(ref_utils)
    gen_net()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
