"""Shared helpers of the GPU parity tests (and of __graft_entry__.smoke)."""
import argparse

import numpy as np


def make_config(**kw):
    base = dict(network='cnn', optimizer='rmsprop', lr=0.00025, reg_param=0.001, momentum=0.9, rmsprop_decay=0.95,
                reward_discount=0.99, minibatch_size=32, max_batch=64, loss='mse', seed=0, input_dtype=None,
                device_index=None, sync=False, task_id=0)
    base.update(kw)
    return argparse.Namespace(**base)


def oracle_params(arch, ishape, A, seed, stddev=0.05):
    from oracle import net_ref as N
    rng = np.random.default_rng(1000 + seed)
    params = N.init_params(arch, ishape, A, seed=seed, stddev=stddev)
    for i in range(1, len(params), 2):
        params[i] += 0.01 * rng.standard_normal(params[i].shape)
    return [p.astype(np.float32).astype(np.float64) for p in params]     # exactly representable in fp32


def make_batch(ishape, A, B, seed, p_term=0.3):
    rng = np.random.default_rng(seed)
    if len(ishape) == 3:
        s = rng.integers(0, 256, (B,) + tuple(ishape), dtype=np.uint8)
        s2 = rng.integers(0, 256, (B,) + tuple(ishape), dtype=np.uint8)
    else:
        s = rng.standard_normal((B,) + tuple(ishape))
        s2 = rng.standard_normal((B,) + tuple(ishape))
    a = rng.integers(0, A, B).astype(np.int32)
    r = rng.integers(-1, 2, B).astype(np.float64) + (rng.random(B) < 0.2) * 0.37
    t = (rng.random(B) < p_term).astype(np.uint8)
    return s, a, r, s2, t


def to_device_batch(net, s, a, r, s2, t):
    import torch
    from replay_memory import Batch
    dev = net.device
    return Batch(net._state_tensor(s), net._state_tensor(s2), torch.as_tensor(a, device=dev),
                 torch.as_tensor(r, device=dev), torch.as_tensor(t, device=dev), None)


def rel_err(got, ref):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    scale = max(np.abs(ref).max(), 1e-30)
    return float(np.abs(got - ref).max() / scale)


def smoke_check():
    """One tiny update of the reference conv net on cuda:0 against the float64 oracle."""
    from oracle import net_ref as N
    import network
    arch, ishape, A, B = 'cnn', (84, 84, 4), 6, 4
    cfg = make_config(network=arch, minibatch_size=B)
    net = network.Network.create_network(cfg, ishape, A)
    params = oracle_params(arch, ishape, A, 0)
    net.set_params(params)
    net.sync_target()
    s, a, r, s2, t = make_batch(ishape, A, B, 0)
    ref = N.RefTrainer(arch, ishape, A, optimizer='rmsprop', lr=cfg.lr, reg_param=cfg.reg_param, gamma=cfg.reward_discount,
                       params=params)
    loss_ref, q_ref, _, _ = ref.train(s, a, r, s2, t)
    loss, q = net.train_on_batch(to_device_batch(net, s, a, r, s2, t), want_loss=True, want_q=True)
    assert rel_err(q.cpu().numpy(), q_ref) < 1e-4, rel_err(q.cpu().numpy(), q_ref)
    assert abs(float(loss.item()) - loss_ref) < 1e-4 * abs(loss_ref)
    for got, exp in zip(net.get_params(), ref.params):
        assert rel_err(got, exp) < 1e-4
    smoke_check_bf16_from_ring()


def smoke_check_bf16_from_ring():
    """The path bench.py times, in small: transitions go into the device frame ring, one sample-to-update call
    (gather + tcgen05 forward/backward + optimizer) in the bf16 tensor-core mode, checked against the oracle with the
    bf16 rounding restated (2e-2, north_star's bf16 tolerance); then a large gather checked bit-exactly."""
    import torch
    from oracle import net_ref as N
    import network
    import replay_memory
    arch, ishape, A, B = 'cnn', (84, 84, 4), 6, 8
    cfg = make_config(network=arch, minibatch_size=B, precision='bf16')
    net = network.Network.create_network(cfg, ishape, A)
    params = oracle_params(arch, ishape, A, 7)
    net.set_params(params)
    net.sync_target()
    s, a, r, s2, t = make_batch(ishape, A, B, 8)
    rm = replay_memory.ReplayMemory(64)
    for i in range(B):
        rm.add(s[i], int(a[i]), float(r[i]), s2[i], bool(t[i]))
    ref = N.RefTrainer(arch, ishape, A, optimizer='rmsprop', lr=cfg.lr, reg_param=cfg.reg_param, gamma=cfg.reward_discount,
                       params=params, bf16=True)
    loss_ref, q_ref, _, _ = ref.train(s, a, r, s2, t)
    idx = torch.arange(B, dtype=torch.int32, device=net.device)
    loss = net.train_step(rm, B, d_indices=idx, want_loss=True)
    assert abs(float(loss.item()) - loss_ref) < 2e-2 * abs(loss_ref), (float(loss.item()), loss_ref)
    for got, exp in zip(net.get_params(), ref.params):
        assert rel_err(got, exp) < 3e-2, rel_err(got, exp)          # one sign-like RMSProp step (see test_net_bf16_gpu.py)
    # gather at a bandwidth-bound size: frames appended on the device, 8192 chained transitions, checked by indexing
    cap, k = 9000, 8192
    big = replay_memory.ReplayMemory(cap, device=net.device.index, trusted=True)
    big._ensure(s[0])
    frames = torch.randint(0, 256, (cap + 4, 84 * 84), dtype=torch.uint8, device=net.device)
    import _cabi as c
    pos = c.c_i64()
    c.check(c.lib.dqn_replay_append_frames_device(big.handle, frames.data_ptr(), cap + 4, c.C.byref(pos), c.current_stream()))
    for i in range(cap):
        c.lib.dqn_replay_add(big.handle, pos.value + 3 + i, pos.value + 4 + i, i % A, float(i % 3 - 1), 0)
    big._resync()
    perm = torch.randperm(cap, device=net.device)[:k].to(torch.int32)
    b = big.gather(perm)
    torch.cuda.synchronize()
    pl = perm.long()
    want = torch.stack([frames[pl + j] for j in range(4)], dim=-1).view(k, 84, 84, 4)
    want_next = torch.stack([frames[pl + 1 + j] for j in range(4)], dim=-1).view(k, 84, 84, 4)
    assert torch.equal(b.states, want) and torch.equal(b.next_states, want_next)
    assert torch.equal(b.actions.long(), pl % A)
