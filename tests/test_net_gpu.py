"""GPU parity of the Q-network kernels (through the C ABI) against the float64 oracle.
Tolerance: 1e-4 relative (north_star's fp32 mode), measured per tensor as max|diff| / max|ref|."""
import numpy as np
import pytest

from gpu_helpers import make_batch, make_config, oracle_params, rel_err, to_device_batch

pytestmark = pytest.mark.gpu
TOL = 1e-4

NETS = [("cnn", (84, 84, 4), 6, 5), ("cnn", (84, 84, 4), 4, 37), ("nature", (84, 84, 4), 6, 3),
        ("simple", (4,), 2, 7), ("simple", (6,), 3, 100)]


def _net(arch, ishape, A, B, **kw):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import network
    cfg = make_config(network=arch, minibatch_size=B, max_batch=max(B, 8), **kw)
    net = network.Network.create_network(cfg, ishape, A)
    return net, cfg


@pytest.mark.parametrize("arch,ishape,A,B", NETS)
def test_forward_q_values_and_argmax(arch, ishape, A, B):
    from oracle import net_ref as N
    net, cfg = _net(arch, ishape, A, B)
    params = oracle_params(arch, ishape, A, 1)
    net.set_params(params)
    tparams = [p * 0.5 for p in params]
    net.set_params(tparams, which='target')
    s, _, _, _, _ = make_batch(ishape, A, B, 2)
    q, am = net.predict(s, argmax=True)
    q_ref = N.forward(arch, params, s, ishape)
    assert rel_err(q.cpu().numpy(), q_ref) < TOL
    assert np.array_equal(am.cpu().numpy(), q.cpu().numpy().argmax(axis=1))
    qt = net.predict(s, use_target_network=True)
    assert rel_err(qt.cpu().numpy(), N.forward(arch, tparams, s, ishape)) < TOL


@pytest.mark.parametrize("arch,ishape,A,B", NETS)
def test_loss_and_every_gradient_tensor(arch, ishape, A, B):
    from oracle import net_ref as N
    net, cfg = _net(arch, ishape, A, B)
    params = oracle_params(arch, ishape, A, 3)
    net.set_params(params)
    tparams = [p + 0.003 * np.random.default_rng(9).standard_normal(p.shape) for p in params]
    tparams = [p.astype(np.float32).astype(np.float64) for p in tparams]
    net.set_params(tparams, which='target')
    s, a, r, s2, t = make_batch(ishape, A, B, 4)
    y = N.td_targets(r, t, N.forward(arch, tparams, s2, ishape), cfg.reward_discount)
    loss_ref, q_ref, g_ref = N.loss_and_grads(arch, params, s, y.astype(np.float64), a, cfg.reg_param, ishape)
    loss, q = net.backward_on_batch(to_device_batch(net, s, a, r, s2, t), want_loss=True, want_q=True, include_reg=True)
    assert rel_err(q.cpu().numpy(), q_ref) < TOL
    assert abs(float(loss.item()) - loss_ref) <= TOL * abs(loss_ref)
    grads = net.get_grads()
    for i, (g, e) in enumerate(zip(grads, g_ref)):
        assert g.shape == e.shape
        assert rel_err(g, e) < TOL, (i, rel_err(g, e))
    # data term only (what the data-parallel all-reduce carries)
    net.backward_on_batch(to_device_batch(net, s, a, r, s2, t), include_reg=False)
    g_noreg = net.get_grads()
    for i in N.reg_indices(arch):
        assert rel_err(g_noreg[i], g_ref[i] - cfg.reg_param * params[i]) < 5 * TOL


ALL_OPTS = ["sgd", "momentum", "rmsprop", "adam", "adagrad", "adadelta", "ftrl"]
# End-to-end updates on the conv net use the four optimizers whose update is well conditioned in fp32 (the two
# presets + sgd/momentum).  adagrad/adadelta/ftrl divide by sqrt(accumulator) (ftrl also subtracts nearly equal
# terms), which amplifies the ~1e-7*max|g| rounding of small gradient entries beyond 1e-4 on unscaled 0..255
# inputs -- TF's own fp32 kernels would differ from a float64 oracle the same way.  Their update RULES are
# checked exactly in test_update_rules_from_a_given_gradient, and end to end on the small net.
STEP_CASES = [(NETS[0], o) for o in ALL_OPTS[:4]] + [(NETS[3], o) for o in ALL_OPTS]


@pytest.mark.parametrize("net_case,opt", STEP_CASES)
def test_three_updates_match_reference_trainer(net_case, opt):
    from oracle import net_ref as N
    arch, ishape, A, B = net_case
    lr = {"sgd": 0.001, "momentum": 0.001, "rmsprop": 0.00025, "adam": 0.001, "adagrad": 0.01, "adadelta": 1.0,
          "ftrl": 0.01}[opt]
    if arch != "simple" and opt in ("sgd", "momentum"):
        lr = 1e-7          # unscaled 0..255 inputs make raw gradients O(1e3); keep plain SGD stable
    net, cfg = _net(arch, ishape, A, B, optimizer=opt, lr=lr)
    params = oracle_params(arch, ishape, A, 5)
    net.set_params(params)
    net.sync_target()
    ref = N.RefTrainer(arch, ishape, A, optimizer=opt, lr=lr, reg_param=cfg.reg_param, gamma=cfg.reward_discount,
                       momentum=cfg.momentum, rmsprop_decay=cfg.rmsprop_decay, target_update_freq=2, params=params)
    for step in range(3):
        s, a, r, s2, t = make_batch(ishape, A, B, 10 + step)
        loss_ref, q_ref, _, _ = ref.train(s, a, r, s2, t)
        loss, q = net.train_on_batch(to_device_batch(net, s, a, r, s2, t), want_loss=True, want_q=True)
        if (step + 1) % 2 == 0:
            net.sync_target()                                   # dqn_agent.py:143, 220
        assert rel_err(q.cpu().numpy(), q_ref) < 2 * TOL, (step, rel_err(q.cpu().numpy(), q_ref))
        assert abs(float(loss.item()) - loss_ref) <= 2 * TOL * abs(loss_ref)
    assert net.get_global_step() == 3
    wtol = TOL
    for i, (g, e) in enumerate(zip(net.get_params(), ref.params)):
        assert rel_err(g, e) < wtol, (opt, i, rel_err(g, e))
    for i, (g, e) in enumerate(zip(net.get_params('target'), ref.target)):
        assert rel_err(g, e) < TOL
    if ref.slots:
        for i, (g, e) in enumerate(zip(net.get_params('slot0'), ref.slots[0])):
            assert rel_err(g, e) < 10 * wtol, (opt, 'slot0', i, rel_err(g, e))


@pytest.mark.parametrize("opt", ALL_OPTS)
def test_update_rules_from_a_given_gradient(opt):
    """The fused optimizer kernel alone: upload a gradient, apply twice, compare weights and slots with the
    TF update rules restated in the oracle (network.py:169-202)."""
    from oracle import net_ref as N
    arch, ishape, A, B = NETS[0]
    lr = {"sgd": 0.01, "momentum": 0.01, "rmsprop": 0.00025, "adam": 0.001, "adagrad": 0.01, "adadelta": 1.0,
          "ftrl": 0.01}[opt]
    net, cfg = _net(arch, ishape, A, B, optimizer=opt, lr=lr)
    params = oracle_params(arch, ishape, A, 12)
    net.set_params(params)
    ref_p = [p.astype(np.float32) for p in params]                 # float32 like TF's own Apply* kernels
    slots = N.init_slots(opt, ref_p)
    rng = np.random.default_rng(3)
    for step in (1, 2):
        s, a, r, s2, t = make_batch(ishape, A, B, 20 + step)
        net.backward_on_batch(to_device_batch(net, s, a, r, s2, t))  # bumps global_step (Adam's t)
        g = [(rng.standard_normal(p.shape) * (10.0 ** rng.integers(-3, 2))).astype(np.float32) for p in ref_p]
        net.set_params(g, which='grad')
        net.apply_gradients()
        full = [gi + np.float32(cfg.reg_param) * p if i in N.reg_indices(arch) else gi
                for i, (gi, p) in enumerate(zip(g, ref_p))]
        N.apply_optimizer(opt, ref_p, full, slots, step, np.float32(lr) if opt != "adam" else lr,
                          np.float32(cfg.momentum), np.float32(cfg.rmsprop_decay))
    tol = 2e-4 if opt == "ftrl" else 2e-5       # ftrl subtracts nearly equal terms
    for i, (got, exp) in enumerate(zip(net.get_params(), ref_p)):
        assert rel_err(got, exp) < tol, (opt, i, rel_err(got, exp))
    for si, sl in enumerate(slots):
        for i, (got, exp) in enumerate(zip(net.get_params('slot%d' % si), sl)):
            assert rel_err(got, exp) < tol, (opt, 'slot', si, i, rel_err(got, exp))


def test_huber_variant_and_all_terminal_batch():
    from oracle import net_ref as N
    arch, ishape, A, B = "cnn", (84, 84, 4), 6, 6
    net, cfg = _net(arch, ishape, A, B, loss='huber', optimizer='sgd', lr=0.01)
    params = oracle_params(arch, ishape, A, 6)
    net.set_params(params)
    net.sync_target()
    s, a, r, s2, t = make_batch(ishape, A, B, 7, p_term=1.1)     # every transition terminal: y = r
    r = r * 3.0
    assert t.all()
    loss, q = net.backward_on_batch(to_device_batch(net, s, a, r, s2, t), want_loss=True, want_q=True, include_reg=True)
    q_ref = N.forward(arch, params, s, ishape)
    diff = q_ref[np.arange(B), a] - r.astype(np.float32)
    hub = np.where(np.abs(diff) <= 1, 0.5 * diff ** 2, np.abs(diff) - 0.5).mean()
    reg = sum(0.5 * np.sum(params[i] ** 2) for i in N.reg_indices(arch))
    assert abs(float(loss.item()) - (hub + cfg.reg_param * reg)) < TOL * abs(hub + cfg.reg_param * reg)
    # bias gradient of the head: clipped error / B on the taken action
    gb = net.get_grads()[9]
    exp = np.zeros(A)
    np.add.at(exp, a, np.clip(diff, -1, 1) / B)
    assert rel_err(gb, exp) < TOL


def test_session_run_compat_path_matches_oracle():
    """The reference's own call pattern: session.run(q_output / train_op, feed_dict) (dqn_agent.py:132-138, 189)."""
    from oracle import net_ref as N
    import network
    arch, ishape, A, B = "simple", (4,), 2, 9
    net, cfg = _net(arch, ishape, A, B, optimizer='adam', lr=0.001, reward_discount=0.9)
    params = oracle_params(arch, ishape, A, 8)
    net.set_params(params)
    sess = network.Session()
    sess.run(net.target_update_ops)
    s, a, r, s2, t = make_batch(ishape, A, B, 9)
    q_next = sess.run(net.target_q_output, feed_dict={net.x_placeholder: list(s2)})
    assert rel_err(q_next, N.forward(arch, params, s2, ishape)) < TOL
    expected_q = [float(r[i]) + 0.9 * (0.0 if t[i] else float(q_next[i].max())) for i in range(B)]
    onehots = [np.eye(A)[a[i]] for i in range(B)]
    _, loss = sess.run([net.train_op, net.summary_op], feed_dict={net.x_placeholder: list(s), net.q_placeholder: expected_q,
                                                                   net.action_placeholder: onehots})
    ref = N.RefTrainer(arch, ishape, A, optimizer='adam', lr=0.001, reg_param=cfg.reg_param, gamma=0.9, params=params)
    loss_ref, _, _, _ = ref.train(s, a, r, s2, t)
    assert abs(loss - loss_ref) <= 2 * TOL * abs(loss_ref)
    for g, e in zip(net.get_params(), ref.params):
        assert rel_err(g, e) < TOL
    assert sess.run(net.global_step) == 1


def test_unsupported_configurations_raise_like_the_reference():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import network
    with pytest.raises(RuntimeError, match="Unsupported network type"):
        network.Network.create_network(make_config(network='resnet'), (84, 84, 4), 6)
    with pytest.raises(RuntimeError, match="Unsupported optimizer"):
        network.Network.create_network(make_config(network='cnn', optimizer='lion'), (84, 84, 4), 6)
    with pytest.raises(RuntimeError, match="1-d input"):
        network.Network.create_network(make_config(network='simple'), (84, 84, 4), 2)
    with pytest.raises(RuntimeError, match="3-d input"):
        network.Network.create_network(make_config(network='cnn'), (4,), 2)


def test_sample_to_update_call_equals_gather_then_train():
    """dqn_train_step (ring -> update in one C call) == gather + train_on_batch, bit for bit."""
    import random
    import torch
    import network
    import replay_memory
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    A, B = 6, 16
    rng = np.random.default_rng(0)
    rm = replay_memory.ReplayMemory(200)
    state = np.repeat(rng.integers(0, 256, (84, 84, 1), dtype=np.uint8), 4, axis=2)
    for i in range(120):
        nxt = np.concatenate([state[..., 1:], rng.integers(0, 256, (84, 84, 1), dtype=np.uint8)], axis=-1)
        rm.add(state, int(rng.integers(0, A)), float(rng.integers(-1, 2)), nxt, bool(i % 17 == 16))
        state = nxt
    nets = []
    for _ in range(2):
        net, cfg = _net("cnn", (84, 84, 4), A, B)
        net.set_params(oracle_params("cnn", (84, 84, 4), A, 11))
        net.sync_target()
        nets.append(net)
    random.seed(3)
    idx = rm.sample_indices(B)
    d_idx = rm.upload_indices(idx)
    l1 = nets[0].train_step(rm, B, d_indices=d_idx, want_loss=True).item()
    batch = rm.gather(d_idx)
    l2, _ = nets[1].train_on_batch(batch, want_loss=True)
    assert l1 == l2.item()
    for g, e in zip(nets[0].get_params(), nets[1].get_params()):
        assert np.array_equal(g, e)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_update_on_an_index_table_equals_update_on_the_same_indices(precision):
    """dqn_train_step_table: three updates that read their rows of a device-resident index table (and advance its
    cursor themselves, also inside one captured graph) == three dqn_train_step calls on the same rows, bit for bit.
    bf16: the conv1 kernels read the table directly; fp32: a small kernel copies the row first."""
    import random
    import torch
    import replay_memory
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    A, B = 6, 16
    rng = np.random.default_rng(0)
    rm = replay_memory.ReplayMemory(200)
    state = np.repeat(rng.integers(0, 256, (84, 84, 1), dtype=np.uint8), 4, axis=2)
    for i in range(120):
        nxt = np.concatenate([state[..., 1:], rng.integers(0, 256, (84, 84, 1), dtype=np.uint8)], axis=-1)
        rm.add(state, int(rng.integers(0, A)), float(rng.integers(-1, 2)), nxt, bool(i % 17 == 16))
        state = nxt
    nets = []
    for _ in range(2):
        net, cfg = _net("cnn", (84, 84, 4), A, B, precision=precision)
        net.set_params(oracle_params("cnn", (84, 84, 4), A, 11))
        net.sync_target()
        nets.append(net)
    random.seed(3)
    rows = [rm.sample_indices(B) for _ in range(2)]          # 2 rows, 3 updates: the third wraps to row 0
    table = replay_memory.IndexTable(rows, torch.device("cuda", 0))
    losses = []
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        losses.append(nets[0].train_step_table(rm, table, want_loss=True).item())
        g = torch.cuda.CUDAGraph()
        loss_buf = torch.zeros(1, dtype=torch.float32, device="cuda")
        with torch.cuda.graph(g, stream=side):
            nets[0].train_step_table(rm, table, want_loss=True, loss_out=loss_buf)
        torch.cuda.synchronize()
        assert int(table.cursor.item()) == 1             # capture does not execute
        for _ in range(2):
            g.replay()
            losses.append(loss_buf.item())
    assert int(table.cursor.item()) == 3
    for u in range(3):
        l = nets[1].train_step(rm, B, d_indices=rm.upload_indices(rows[u % 2]), want_loss=True).item()
        assert l == losses[u], (u, l, losses[u])
    for a, e in zip(nets[0].get_params(), nets[1].get_params()):
        assert np.array_equal(a, e)
