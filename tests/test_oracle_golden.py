"""CPU: the oracle against the fixtures recorded from the reference's own code (data half), and the two
independent restatements of the TF graph against each other + finite differences (model half, unpinned)."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import net_ref as N
from oracle import torch_ref as T
from oracle.gen_golden import episode_script
from oracle.replay_ref import (FrameBufferRef, ReplayRef, decay_per_step_ref, one_hot_ref, partition_ref)


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


def test_sample_index_stream_matches_reference(golden_dir):
    g = _load(golden_dir, "replay_sample_indices.npz")
    for ci in range(int(g["ncases"][0])):
        adds, cap, k, seed = (int(v) for v in g["case%d_cfg" % ci])
        rm = ReplayRef(cap)
        for i in range(adds):
            rm.add(i, i % 6, float(i % 3 - 1), i + 1, (i % 11) == 10)
        assert [rm.size(), rm.capacity()] == list(g["case%d_size" % ci])
        random.seed(seed)
        nt, t = rm.get_minibatch(k)
        if "case%d_none" % ci in g.files:
            assert nt is None and t is None
            continue
        nt, t = list(nt), list(t)
        assert [x[0] for x in nt] == list(g["case%d_nt" % ci])
        assert [x[0] for x in t] == list(g["case%d_t" % ci])
        assert list(ReplayRef.get_next_states(nt)) == list(g["case%d_next_nt" % ci])
        assert list(ReplayRef.get_states(t)) == list(g["case%d_state_t" % ci])


def test_framebuffer_matches_reference(golden_dir):
    g = _load(golden_dir, "framebuffer.npz")
    fr = g["frames"]
    fb = FrameBufferRef(4)
    assert fb.get_state() is None
    for i in range(6):
        fb.append(fr[i])
        assert np.array_equal(fb.get_state(), g["fb4_after%d" % i])
    fb.clear()
    assert fb.get_state() is None and bool(g["fb4_cleared_is_none"][0])
    fb.append(fr[6])
    st = fb.get_state()
    assert np.array_equal(st, g["fb4_restart"])
    assert [st.shape[0], st.shape[1], st.shape[2], st.strides[0], st.strides[1], st.strides[2],
            int(st.flags["C_CONTIGUOUS"])] == list(g["fb4_meta"])
    fb1 = FrameBufferRef(1)
    fb1.append(fr[0]); fb1.append(fr[1])
    assert np.array_equal(fb1.get_state(), g["fb1_state"])
    fb2 = FrameBufferRef(2, preprocessor=lambda f: (f // 2).astype(np.uint8))
    fb2.append(fr[2]); fb2.append(fr[3]); fb2.append(fr[4])
    assert np.array_equal(fb2.get_state(), g["fb2_pre_state"])
    with pytest.raises(RuntimeError) as e:
        FrameBufferRef(0)
    assert int(g["fb0_raises"][0]) == 1 and str(e.value) == str(g["fb0_msg"])


def test_product_framebuffer_matches_reference_fixture(golden_dir):
    """The shipped FrameBuffer (dist-dqn_b200/frame_buffer.py, pure host code) against what the reference's own class
    returned (frame_buffer.py:27-57), plus the property the agent's shortcut relies on: ``frames[-1]`` is the newest
    frame, i.e. the last channel of ``get_state()``."""
    from frame_buffer import FrameBuffer
    g = _load(golden_dir, "framebuffer.npz")
    fr = g["frames"]
    fb = FrameBuffer(4)
    assert fb.get_state() is None
    for i in range(6):
        fb.append(fr[i])
        st = fb.get_state()
        assert np.array_equal(st, g["fb4_after%d" % i])
        assert np.array_equal(fb.frames[-1], st[..., -1]) and fb.frames[-1] is not None
    fb.clear()
    assert fb.get_state() is None
    fb.append(fr[6])
    st = fb.get_state()
    assert np.array_equal(st, g["fb4_restart"])
    assert [st.shape[0], st.shape[1], st.shape[2], st.strides[0], st.strides[1], st.strides[2],
            int(st.flags["C_CONTIGUOUS"])] == list(g["fb4_meta"])
    fb1 = FrameBuffer(1)
    fb1.append(fr[0]); fb1.append(fr[1])
    assert np.array_equal(fb1.get_state(), g["fb1_state"])
    fb2 = FrameBuffer(2, preprocessor=lambda f: (f // 2).astype(np.uint8))
    fb2.append(fr[2]); fb2.append(fr[3]); fb2.append(fr[4])
    assert np.array_equal(fb2.get_state(), g["fb2_pre_state"])
    with pytest.raises(RuntimeError) as e:
        FrameBuffer(0)
    assert str(e.value) == str(g["fb0_msg"])


@pytest.mark.parametrize("name", ["small", "atari", "vec"])
def test_episode_script_matches_reference(golden_dir, name):
    g = _load(golden_dir, "episode_script_%s.npz" % name)
    batches, size = episode_script(ReplayRef, FrameBufferRef, g["frames"], list(g["ep_lens"]), list(g["ep_done"]),
                                   int(g["capacity"][0]), int(g["k"][0]), int(g["sample_every"][0]),
                                   int(g["seed"][0]), int(g["frames_per_state"][0]))
    assert size == int(g["final_size"][0]) and len(batches) == int(g["nbatches"][0]) and len(batches) > 3
    for i, b in enumerate(batches):
        for key, v in b.items():
            assert np.array_equal(np.asarray(v), g["b%d_%s" % (i, key)]), (i, key)


def test_utils_match_reference(golden_dir):
    g = _load(golden_dir, "utils.npz")
    assert np.array_equal(one_hot_ref(2, 6), g["one_hot_2_6"]) and one_hot_ref(2, 6).dtype == np.float64
    assert np.array_equal(one_hot_ref(0, 2), g["one_hot_0_2"])
    got = [decay_per_step_ref(1.0, 0.1, 1000000), decay_per_step_ref(0.5, 0.1, 50000),
           decay_per_step_ref(1.0, 0.1, 0), decay_per_step_ref(1.0, 0.1, -3)]
    assert got == list(g["decay_per_step"])
    a, b = partition_ref(lambda v: v % 3 == 0, [5, 3, 9, 1, 6, 7, 0, 2])
    assert list(a) == list(g["partition_false"]) and list(b) == list(g["partition_true"])


# ---------------------------------------------------------------- model half (unpinned: self-consistency)
CASES = [("cnn", (84, 84, 4), 6, 3), ("nature", (84, 84, 4), 4, 2), ("simple", (4,), 2, 5)]


def _batch(arch, ishape, A, B, seed):
    rng = np.random.default_rng(seed)
    params = N.init_params(arch, ishape, A, seed=seed, stddev=0.05)
    for i in range(1, len(params), 2):
        params[i] += 0.01 * rng.standard_normal(params[i].shape)
    x = rng.integers(0, 256, (B,) + ishape, dtype=np.uint8) if len(ishape) == 3 else rng.standard_normal((B,) + ishape)
    y = rng.standard_normal(B)
    a = rng.integers(0, A, B)
    return params, x, y, a


def test_same_padding_table():
    layers, flat = N.conv_plan("cnn", (84, 84, 4))
    assert [(L["oh"], L["ph"], L["pad_h"]) for L in layers] == [(21, 11, (2, 2)), (6, 3, (1, 2)), (3, 2, (1, 1))]
    assert flat == 256
    assert sum(int(np.prod(s)) for s in N.param_shapes("cnn", (84, 84, 4), 6)) == 145318
    assert sum(int(np.prod(s)) for s in N.param_shapes("cnn", (84, 84, 4), 4)) == 144804
    assert sum(int(np.prod(s)) for s in N.param_shapes("simple", (4,), 2)) == 562
    layers, flat = N.conv_plan("nature", (84, 84, 4))
    assert [L["oh"] for L in layers] == [20, 9, 7] and flat == 3136
    assert N.flops_per_sample("cnn", (84, 84, 4), 6) == (5191168, 17152000)


@pytest.mark.parametrize("arch,ishape,A,B", CASES)
def test_numpy_and_torch_restatements_agree(arch, ishape, A, B):
    params, x, y, a = _batch(arch, ishape, A, B, 7)
    loss, q, grads = N.loss_and_grads(arch, params, x, y, a, 0.001, ishape)
    tp = [torch.tensor(p, dtype=torch.float64) for p in params]
    tl, tq, tg = T.loss_and_grads(arch, tp, torch.tensor(x), torch.tensor(y), torch.tensor(a), 0.001, ishape)
    assert abs(float(tl) - loss) <= 1e-10 * max(1.0, abs(loss))
    assert np.allclose(tq.numpy(), q, rtol=1e-10, atol=1e-10)
    for g, t in zip(grads, tg):
        assert g.shape == tuple(t.shape)
        assert np.allclose(t.numpy(), g, rtol=1e-9, atol=1e-10 * max(1.0, np.abs(g).max()))


@pytest.mark.parametrize("arch,ishape,A,B", [CASES[0], CASES[2]])
def test_finite_difference_gradient(arch, ishape, A, B):
    params, x, y, a = _batch(arch, ishape, A, B, 3)
    x = (x / 64.0) if len(ishape) == 3 else x           # keep activations O(1) so FD is well conditioned
    loss, _, grads = N.loss_and_grads(arch, params, x, y, a, 0.001, ishape)
    rng = np.random.default_rng(0)
    for pi in range(len(params)):
        for _ in range(3):
            idx = tuple(rng.integers(0, s) for s in params[pi].shape)
            eps = 1e-6
            old = params[pi][idx]
            params[pi][idx] = old + eps
            lp = N.loss_and_grads(arch, params, x, y, a, 0.001, ishape)[0]
            params[pi][idx] = old - eps
            lm = N.loss_and_grads(arch, params, x, y, a, 0.001, ishape)[0]
            params[pi][idx] = old
            fd = (lp - lm) / (2 * eps)
            assert abs(fd - grads[pi][idx]) <= 1e-5 * max(1.0, abs(fd)), (pi, idx, fd, grads[pi][idx])


@pytest.mark.parametrize("opt", N.OPTIMIZERS)
def test_optimizer_rules_first_step_closed_form(opt):
    w = [np.array([1.0, -2.0, 0.5])]
    g = [np.array([0.1, -0.2, 0.0])]
    slots = N.init_slots(opt, w)
    N.apply_optimizer(opt, w, g, slots, 1, lr=0.01, momentum=0.9, rmsprop_decay=0.95)
    w0, gg = np.array([1.0, -2.0, 0.5]), g[0]
    exp = {
        "sgd": w0 - 0.01 * gg,
        "momentum": w0 - 0.01 * gg,
        "rmsprop": w0 - 0.01 * gg / np.sqrt(1 + 0.05 * (gg * gg - 1) + 1e-10),
        "adam": w0 - 0.01 * np.sqrt(1 - 0.999) / (1 - 0.9) * (0.1 * gg) / (np.sqrt(0.001 * gg * gg) + 1e-8),
        "adagrad": w0 - 0.01 * gg / np.sqrt(0.1 + gg * gg),
        "adadelta": w0 - 0.01 * np.sqrt(1e-8) / np.sqrt(0.05 * gg * gg + 1e-8) * gg,
        "ftrl": -(gg - (np.sqrt(0.1 + gg * gg) - np.sqrt(0.1)) / 0.01 * w0) / (np.sqrt(0.1 + gg * gg) / 0.01),
    }[opt]
    assert np.allclose(w[0], exp, rtol=1e-12, atol=1e-15)


def test_td_targets_follow_reference_rule():
    q = np.array([[1.0, 3.0], [2.0, -1.0], [0.5, 0.25]], np.float32)
    y = N.td_targets([1.0, -1.0, 0.0], [False, True, False], q, 0.99)
    assert y.dtype == np.float32
    assert np.array_equal(y, np.array([1.0 + 0.99 * 3.0, -1.0, 0.99 * 0.5]).astype(np.float32))


def test_net_regression_fixture(golden_dir):
    g = _load(golden_dir, "net_ref_regression.npz")
    from oracle.gen_golden import _sample
    ci = 0
    while "c%d_cfg" % ci in g.files:
        arch, opt, A, B, lr, gamma, ishape = (str(v) for v in g["c%d_cfg" % ci])
        A, B, lr, gamma, ishape = int(A), int(B), float(lr), float(gamma), eval(ishape)
        rng = np.random.default_rng(100 + ci)
        tr = N.RefTrainer(arch, ishape, A, optimizer=opt, lr=lr, reg_param=0.001, gamma=gamma, seed=ci,
                          params=N.init_params(arch, ishape, A, seed=ci, stddev=0.05), target_update_freq=2)
        for i in range(1, len(tr.params), 2):
            tr.params[i] += 0.01 * rng.standard_normal(tr.params[i].shape)
        tr.target = [p + 0.002 * rng.standard_normal(p.shape) for p in tr.params]
        mk = (lambda: rng.integers(0, 256, (B,) + ishape, dtype=np.uint8)) if len(ishape) == 3 else \
            (lambda: rng.standard_normal((B,) + ishape))
        for step in range(3):
            s, s2 = mk(), mk()
            a = rng.integers(0, A, B)
            r = rng.integers(-1, 2, B).astype(np.float64)
            term = rng.random(B) < 0.3
            loss, q, grads, y = tr.train(s, a, r, s2, term)
            assert np.allclose(loss, g["c%d_s%d_loss" % (ci, step)][0], rtol=1e-9)
            assert np.allclose(q, g["c%d_s%d_q" % (ci, step)], rtol=1e-9, atol=1e-12)
            assert np.array_equal(y, g["c%d_s%d_y" % (ci, step)])
            for pi in range(len(grads)):
                assert np.allclose(_sample(grads[pi]), g["c%d_s%d_g%d" % (ci, step, pi)], rtol=1e-7, atol=1e-12)
                assert np.allclose(_sample(tr.params[pi]), g["c%d_s%d_p%d" % (ci, step, pi)], rtol=1e-9, atol=1e-12)
        ci += 1
    assert ci == 5
