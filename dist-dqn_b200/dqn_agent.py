"""DQNAgent with the reference's surface (reference src/dqn_agent.py:13-285).

The episode loop, epsilon schedule, replay prefill and target-sync cadence follow the reference;
what changes is ``_train_minibatch``: the reference makes three ``session.run`` round trips with
two host->device uploads of float32 frame stacks and a Python loop for the TD targets
(dqn_agent.py:108-143, 224-253); here the host only draws the sample indices (same global
``random`` stream as the reference, so the same transitions are picked for the same seed) and
the rest is one CUDA-graph launch: gather -> target/online forward -> TD target -> loss ->
backward -> optimizer.  With more than one rank the gradient is averaged across GPUs with an NCCL
all-reduce inside the same graph (the reference's --sync semantics, network.py:187-192).
"""
from __future__ import annotations

import json
import logging
import os
import random
from functools import partial

import numpy as np

import utils
from frame_buffer import FrameBuffer
from stats import Stats


class SummaryWriter(object):
    """Scalar summaries as JSON lines under ``logdir`` (stands in for tf.train.SummaryWriter,
    dqn_agent.py:39; only the 'loss' scalar exists in the reference, network.py:154-155)."""

    def __init__(self, logdir):
        os.makedirs(logdir, exist_ok=True)
        self._f = open(os.path.join(logdir, 'summaries.jsonl'), 'a')

    def add_summary(self, value, step):
        self._f.write(json.dumps({'step': int(step), 'loss': float(value)}) + '\n')
        self._f.flush()


class DQNAgent(object):
    # Reward penalty on failure for each environment (dqn_agent.py:15-17; empty in the reference)
    FAILURE_PENALTY = {}

    def __init__(self, env, network, session, replay_memory, config, enable_summary=True):
        self.env = env
        self.network = network
        self.session = session                # kept for signature compatibility; the fast path does not use it
        self.replay_memory = replay_memory
        self.config = config

        self.training_steps = 0
        self.stats = Stats()

        self.random_action_prob = config.init_random_action_prob
        self.random_action_prob_decay = utils.decay_per_step(
            init_val=config.init_random_action_prob,
            min_val=config.min_random_action_prob,
            steps=config.random_action_explore_steps,
        )

        self.summary_writer = SummaryWriter(config.logdir) if enable_summary else None
        self.frame_buffer = FrameBuffer(
            frames_per_state=config.frames_per_state,
            preprocessor=self._get_frame_resizer(env, config),
        )
        self._world = _world_size()
        self._graphs = {}
        self.use_cuda_graph = bool(getattr(config, 'cuda_graph', True))
        self.last_loss = None

        self._prefill_replay_memory(config.replay_start_size)     # dqn_agent.py:47
        self._update_target_network()                             # dqn_agent.py:50 (step 0 always syncs)

    # ------------------------------------------------------------------ training loop (dqn_agent.py:52-106)
    def train(self, num_episodes, max_steps_per_episode, supervisor=None):
        for episode in range(num_episodes):
            reward, steps = self.train_episode(max_steps_per_episode)
            self.stats.log_episode(reward, steps)
            logging.info('Episode = %d, steps = %d, reward = %d, training steps = %d, last-100 mean reward = %.2f'
                         % (episode, steps, reward, self.training_steps, self.stats.last_100_mean_reward()))
            if supervisor and supervisor.should_stop():
                logging.warning('Received signal to stop. Exiting train loop.')
                break

    def train_episode(self, max_steps):
        state = self._reset_episode()
        total_reward = steps = 0
        done = False
        while not done and steps < max_steps:
            action = self._pick_action(state)
            observation, reward, done, _ = self.env.step(action)
            total_reward += reward
            steps += 1
            if done:
                reward = self.FAILURE_PENALTY.get(self.env.spec.id, reward)
            state = self._record(state, action, reward, observation, done)
            if steps % self.config.update_freq == 0:
                self._train_minibatch(self.config.minibatch_size)
        return total_reward, steps

    def _reset_episode(self):
        self.frame_buffer.clear()
        self.frame_buffer.append(self.env.reset())
        return self.frame_buffer.get_state()

    def _record(self, state, action, reward, observation, done):
        self.frame_buffer.append(observation)
        next_state = self.frame_buffer.get_state()
        self.replay_memory.add(state, action, reward, next_state, done)
        return next_state

    # ------------------------------------------------------------------ the hot path (dqn_agent.py:108-143)
    def _train_minibatch(self, minibatch_size):
        rm, net = self.replay_memory, self.network
        if rm.size() < minibatch_size:
            return
        d_idx = rm.upload_indices(rm.sample_indices(minibatch_size))     # random.sample on the global generator
        want_loss = self._should_log_summary()
        loss = self._device_step(minibatch_size, d_idx, want_loss)
        if want_loss:
            self.last_loss = float(loss.item())
            self.summary_writer.add_summary(self.last_loss, self.training_steps)
        self.training_steps += 1
        self._update_target_network()

    def _eager_step(self, k, d_idx, want_loss):
        net, rm = self.network, self.replay_memory
        if self._world == 1:
            return net.train_step(rm, k, d_indices=d_idx, want_loss=want_loss)
        if getattr(net, 'peers_connected', False):     # gradient exchange fused into the optimizer (NVLink peer loads)
            return net.dp_train_step(rm, k, d_indices=d_idx, want_loss=want_loss)
        import torch.distributed as dist
        batch = rm.gather(d_idx)
        loss, _ = net.backward_on_batch(batch, want_loss=want_loss)
        dist.all_reduce(net.flat_buffer('grad'))                  # sum; grad_scale = 1/world applied by the update
        net.apply_gradients()
        return loss

    def _device_step(self, k, d_idx, want_loss):
        import torch
        if not self.use_cuda_graph:
            return self._eager_step(k, d_idx, want_loss)
        key = (k, bool(want_loss))
        g = self._graphs.get(key)
        if g is None:
            # first call per shape runs eagerly (loads kernels, sizes buffers), the second is captured
            self._graphs[key] = 'warm'
            return self._eager_step(k, d_idx, want_loss)
        if g == 'warm':
            self.replay_memory.flush()
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = self._eager_step(k, d_idx, want_loss)
            self._graphs[key] = g = (graph, out, d_idx)
        graph, out, captured_idx = g
        if captured_idx.data_ptr() != d_idx.data_ptr():
            captured_idx.copy_(d_idx)
        self.replay_memory.flush()                                # staged frames go in before the graph reads the ring
        graph.replay()
        return out

    # ------------------------------------------------------------------ acting (dqn_agent.py:145-189)
    def _pick_action(self, state):
        if self._roll_random_action_dice():
            return self.env.action_space.sample()
        _, greedy = self.network.predict([state], argmax=True)
        return int(greedy.item())

    def _roll_random_action_dice(self):
        self._decay_random_action_prob()
        return random.random() < self.random_action_prob

    def _decay_random_action_prob(self):
        if self.random_action_prob > self.config.min_random_action_prob:
            self.random_action_prob -= self.random_action_prob_decay

    def _predict_q_values(self, states, use_target_network=False):
        """ndarray [N, A] of Q-values under the online or target parameters."""
        return self.network.predict(states, use_target_network=use_target_network).cpu().numpy()

    # ------------------------------------------------------------------ prefill / target / summaries
    def _prefill_replay_memory(self, prefill_size):
        """Uniform-random policy until the memory holds prefill_size transitions (dqn_agent.py:191-213)."""
        terminal = True
        state = None
        while self.replay_memory.size() < prefill_size:
            if terminal:
                state = self._reset_episode()
            action = self.env.action_space.sample()
            observation, reward, terminal, _ = self.env.step(action)
            state = self._record(state, action, reward, observation, terminal)

    def _update_target_network(self):
        if self.training_steps % self.config.target_update_freq == 0:
            logging.info('Updating target network')
            self.network.sync_target()

    def _should_log_summary(self):
        if self.summary_writer is None:
            return False
        freq = self.config.summary_freq
        return freq > 0 and (self.training_steps % freq == 0)

    @classmethod
    def _get_frame_resizer(cls, env, config):
        width, height = config.resize_width, config.resize_height
        if width > 0 and height > 0:
            return partial(utils.resize_image, width=width, height=height)
        return lambda frame: frame

    @classmethod
    def get_input_shape(cls, env, config):
        width, height = config.resize_width, config.resize_height
        if width > 0 and height > 0:
            return (width, height, config.frames_per_state)
        return env.observation_space.shape


def _world_size():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_world_size()
    except Exception:
        pass
    return 1
