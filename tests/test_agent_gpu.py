"""GPU: agent-level behaviour through the reference's surface -- CartPole with the CONTROL preset (BASELINE config 0),
the Atari-shaped agent loop on synthetic frames, batched greedy acting (config 5), checkpoint/resume."""
import argparse
import os
import random
import tempfile

import numpy as np
import pytest

from gpu_helpers import make_batch, make_config, oracle_params, rel_err, to_device_batch

pytestmark = pytest.mark.gpu


def _cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")


def _control_config(**kw):
    # scripts/dqn_params.sh:7-20 (CONTROL preset), shortened run
    base = dict(network='simple', optimizer='adam', lr=0.001, reg_param=0.001, momentum=0.9, rmsprop_decay=0.95,
                minibatch_size=100, num_episodes=30, max_steps_per_episode=200, replay_memory_capacity=50000,
                target_update_freq=3000, reward_discount=0.9, init_random_action_prob=0.5, min_random_action_prob=0.1,
                random_action_explore_steps=50000, frames_per_state=1, resize_width=0, resize_height=0, update_freq=1,
                replay_start_size=0, summary_freq=10, logdir=tempfile.mkdtemp(prefix='dqn_test_'), loss='mse', seed=0,
                max_batch=100, input_dtype=None, device_index=None, cuda_graph=True)
    base.update(kw)
    return argparse.Namespace(**base)


def test_cartpole_agent_trains_and_matches_oracle_step_by_step():
    """The whole loop on CartPole; every training step is replayed through the oracle with the same sampled
    transitions (the agent draws indices from Python's global generator like the reference) and must agree."""
    _cuda()
    import envs
    from dqn_agent import DQNAgent
    from network import Network, Session
    from replay_memory import ReplayMemory
    from oracle import net_ref as N
    from oracle.replay_ref import ReplayRef
    cfg = _control_config(num_episodes=6, minibatch_size=32, max_batch=32, cuda_graph=False)
    random.seed(7)
    env = envs.make('CartPole-v0', seed=3)
    net = Network.create_network(cfg, DQNAgent.get_input_shape(env, cfg), env.action_space.n)
    ref = N.RefTrainer('simple', (4,), 2, optimizer='adam', lr=cfg.lr, reg_param=cfg.reg_param, gamma=cfg.reward_discount,
                       target_update_freq=cfg.target_update_freq, params=[p.astype(np.float64) for p in net.get_params()])
    rm = ReplayMemory(cfg.replay_memory_capacity)
    agent = DQNAgent(env, net, Session(), rm, cfg, enable_summary=True)
    shadow = ReplayRef(cfg.replay_memory_capacity)
    orig_add, orig_sample = rm.add, rm.sample_indices
    picked = []

    def add(s, a, r, s2, t):
        shadow.add(np.array(s), a, r, np.array(s2), t)
        orig_add(s, a, r, s2, t)

    def sample_indices(k):
        idx = orig_sample(k)
        picked.append(idx)
        return idx
    rm.add, rm.sample_indices = add, sample_indices
    agent.train(cfg.num_episodes, cfg.max_steps_per_episode)
    assert agent.training_steps == len(picked) > 20 and agent.training_steps == net.get_global_step()
    assert rm.size() == shadow.size() == agent.stats.total_steps
    for idx in picked:                                    # replay the same minibatches through the oracle
        items = [shadow.memory[i] for i in idx]
        # the shadow deque only grew afterwards, logical indices are stable because nothing was evicted
        ref.train(np.stack([x[0] for x in items]), [x[1] for x in items], [x[2] for x in items],
                  np.stack([x[3] for x in items]), [x[4] for x in items])
    for g, e in zip(net.get_params(), ref.params):
        assert rel_err(g, e) < 2e-3, rel_err(g, e)        # ~100 Adam steps of fp32 vs float64
    assert agent.last_loss is not None and np.isfinite(agent.last_loss)
    assert os.path.exists(os.path.join(cfg.logdir, 'summaries.jsonl'))


def test_atari_shaped_agent_loop_with_cuda_graph():
    _cuda()
    import envs
    from dqn_agent import DQNAgent
    from network import Network, Session
    from replay_memory import ReplayMemory
    cfg = _control_config(network='cnn', optimizer='rmsprop', lr=0.00025, minibatch_size=32, max_batch=32, num_episodes=2,
                          max_steps_per_episode=60, replay_memory_capacity=500, target_update_freq=20, reward_discount=0.99,
                          frames_per_state=4, resize_width=84, resize_height=84, update_freq=4, replay_start_size=64,
                          summary_freq=5, init_random_action_prob=1.0, precision='bf16')
    random.seed(1)
    env = envs.SyntheticAtariEnv('SyntheticPong-v0', num_actions=6, seed=5, min_len=40, max_len=80)
    net = Network.create_network(cfg, DQNAgent.get_input_shape(env, cfg), env.action_space.n)
    assert DQNAgent.get_input_shape(env, cfg) == (84, 84, 4)
    rm = ReplayMemory(cfg.replay_memory_capacity, trusted=True)
    agent = DQNAgent(env, net, Session(), rm, cfg)
    assert rm.size() >= 64                                # prefilled with the random policy (dqn_agent.py:191-213)
    agent.train(cfg.num_episodes, cfg.max_steps_per_episode)
    assert agent.training_steps >= 20 and net.get_global_step() == agent.training_steps
    assert np.isfinite(agent.last_loss)
    q = agent._predict_q_values([agent.frame_buffer.get_state()])
    assert q.shape == (1, 6) and np.isfinite(q).all()
    assert 0.1 <= agent.random_action_prob < 1.0


def test_graph_embedded_ingest_keeps_the_ring_identical_to_the_reference_deque():
    """The captured step moves the frames staged by ReplayMemory.add into the ring itself (dqn_replay_seal /
    dqn_replay_ingest_staged).  After a graph-driven agent loop -- including episode starts, which stage a full stack --
    the ring must hold exactly what the reference's deque(maxlen) holds (replay_memory.py:19-29), bit for bit."""
    _cuda()
    import collections
    import envs
    from dqn_agent import DQNAgent
    from network import Network, Session
    from replay_memory import ReplayMemory
    cfg = _control_config(network='cnn', optimizer='rmsprop', lr=0.00025, minibatch_size=32, max_batch=32, num_episodes=4,
                          max_steps_per_episode=50, replay_memory_capacity=150, target_update_freq=20, reward_discount=0.99,
                          frames_per_state=4, resize_width=84, resize_height=84, update_freq=4, replay_start_size=40,
                          summary_freq=3, init_random_action_prob=1.0, precision='bf16')
    random.seed(3)
    env = envs.SyntheticAtariEnv('SyntheticPong-v0', num_actions=6, seed=9, min_len=20, max_len=45)
    net = Network.create_network(cfg, DQNAgent.get_input_shape(env, cfg), env.action_space.n)
    rm = ReplayMemory(cfg.replay_memory_capacity, trusted=True)
    mirror = collections.deque(maxlen=cfg.replay_memory_capacity)
    real_add = rm.add

    def add(state, action, reward, observation, terminal, **kw):
        mirror.append((np.array(state, copy=True), int(action), float(reward), np.array(observation, copy=True), bool(terminal)))
        if 'newest' in kw:                                            # the agent's shortcut must be the stack's last frame
            assert np.array_equal(kw['newest'], observation[..., -1])
        real_add(state, action, reward, observation, terminal, **kw)

    rm.add = add
    agent = DQNAgent(env, net, Session(), rm, cfg)
    agent.train(cfg.num_episodes, cfg.max_steps_per_episode)
    assert agent.training_steps >= 30 and len(agent._graphs) >= 1       # the graph path ran (and wrapped the ring: 150 < adds)
    assert rm.size() == len(mirror) == cfg.replay_memory_capacity
    b = rm.sample(rm.size(), indices=list(range(rm.size())))
    states, nexts = b.states.cpu().numpy(), b.next_states.cpu().numpy()
    acts, rews, terms = b.actions.cpu().numpy(), b.rewards.cpu().numpy(), b.terminals.cpu().numpy()
    for i, (s0, a0, r0, s1, t0) in enumerate(mirror):
        assert np.array_equal(states[i], s0) and np.array_equal(nexts[i], s1), i
        assert acts[i] == a0 and rews[i] == r0 and bool(terms[i]) == t0, i


def test_batched_greedy_act_1024_envs():
    """BASELINE config 5: greedy actions for 1024 synthetic env states in one call (chunks of max_batch)."""
    _cuda()
    from oracle import net_ref as N
    import network
    A = 4
    cfg = make_config(network='cnn', minibatch_size=512, max_batch=512)
    net = network.Network.create_network(cfg, (84, 84, 4), A)
    params = oracle_params('cnn', (84, 84, 4), A, 2)
    net.set_params(params)
    s = np.random.default_rng(2).integers(0, 256, (1024, 84, 84, 4), dtype=np.uint8)
    q, am = net.predict(s, argmax=True)
    q_ref = N.forward('cnn', params, s[:64], (84, 84, 4))
    assert rel_err(q[:64].cpu().numpy(), q_ref) < 1e-4
    assert np.array_equal(am.cpu().numpy(), q.cpu().numpy().argmax(axis=1))


def test_checkpoint_round_trip():
    _cuda()
    import network
    cfg = make_config(network='cnn', minibatch_size=8, max_batch=8, optimizer='adam', lr=0.001)
    net = network.Network.create_network(cfg, (84, 84, 4), 6)
    net.set_params(oracle_params('cnn', (84, 84, 4), 6, 4))
    b = make_batch((84, 84, 4), 6, 8, 1)
    net.train_on_batch(to_device_batch(net, *b))
    net.train_on_batch(to_device_batch(net, *b))
    path = os.path.join(tempfile.mkdtemp(prefix='dqn_ckpt_'), 'ckpt.npz')
    net.save(path, extra={'epsilon': 0.37, 'training_steps': 2})
    net.train_on_batch(to_device_batch(net, *b))                       # the step to reproduce after resume
    expect = net.get_params()
    net2 = network.Network.create_network(cfg, (84, 84, 4), 6)
    extra = net2.load(path)
    assert extra == {'epsilon': 0.37, 'training_steps': 2} and net2.get_global_step() == 2
    net2.train_on_batch(to_device_batch(net2, *b))
    for g, e in zip(net2.get_params(), expect):
        assert np.array_equal(g, e)                                     # bit-identical continuation (Adam's t restored)


def test_main_entry_point_runs_the_tensor_core_path():
    """The product entry point with the reference's command line (scripts/dqn_single_node.sh -> main.py): conv net on
    synthetic Atari-shaped frames, bf16 tensor-core mode by default, a few episodes; the summaries file and the
    global_step/sec log line (the reference's throughput signal, main.py:136-143) must appear."""
    _cuda()
    import subprocess
    import sys
    from conftest import ROOT
    logdir = tempfile.mkdtemp(prefix='dqn_main_')
    env = dict(os.environ, DQN_LOG_ROOT=logdir,
               DQN_EXTRA_FLAGS='--num_episodes 3 --max_steps_per_episode 80 --replay_memory_capacity 2000 '
                               '--replay_start_size 64 --summary_freq 4 --seed 3')
    r = subprocess.run(['bash', os.path.join(ROOT, 'scripts', 'dqn_single_node.sh'), 'atari', 'SyntheticPong-v0'],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:]
    assert 'global_step/sec' in r.stdout and 'Episode = 2' in r.stdout, r.stdout[-3000:]
    lines = open(os.path.join(logdir, 'train', 'summaries.jsonl')).read().strip().splitlines()
    assert len(lines) >= 3 and all(np.isfinite(__import__('json').loads(l)['loss']) for l in lines)
