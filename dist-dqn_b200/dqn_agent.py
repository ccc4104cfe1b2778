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

import atexit
import json
import logging
import os
import random
from functools import partial

import time

import numpy as np

import data_parallel
import utils
from frame_buffer import FrameBuffer
from stats import Stats


class SummaryWriter(object):
    """Scalar summaries as JSON lines under ``logdir`` (stands in for tf.train.SummaryWriter,
    dqn_agent.py:39; only the 'loss' scalar exists in the reference, network.py:154-155).  Like the TF writer, which
    queues events for a background thread, ``add_summary`` does not touch the disk: lines are buffered and written out
    every ``flush_every`` summaries, on ``flush()`` and at exit."""

    def __init__(self, logdir, flush_every=256):
        os.makedirs(logdir, exist_ok=True)
        self._f = open(os.path.join(logdir, 'summaries.jsonl'), 'a')
        self._lines = []
        self._flush_every = flush_every
        atexit.register(self.flush)

    def add_summary(self, value, step):
        self._lines.append('{"step": %d, "loss": %r}\n' % (step, float(value)))
        if len(self._lines) >= self._flush_every:
            self.flush()

    def flush(self):
        if self._lines and not self._f.closed:
            self._f.write(''.join(self._lines))
            self._f.flush()
        self._lines = []


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
        # summaries are written one update late: the loss of update s is copied to pinned host memory asynchronously
        # and picked up when update s+1 is launched, so the host never idles on the GPU (flush_summaries() drains it)
        self.async_summary = bool(getattr(config, 'async_summary', True))
        self._pending = None
        self._stage_evt = [None, None, None]
        # our ReplayMemory takes the newest frame as a contiguous array (saves a strided copy per env step); any other
        # object with the reference's 5-argument add() keeps working
        self._pass_newest = (config.frames_per_state > 1 and getattr(replay_memory, '_trusted', False)
                             and type(replay_memory).__name__ == 'ReplayMemory' and hasattr(replay_memory, 'seal'))

        # data parallel: the number of gradient exchanges must not depend on the rank.  The "replay memory still
        # smaller than a minibatch" early-out of _train_minibatch would make it so, hence the prefill covers at least
        # one minibatch; ranks whose episode loop ends early keep updating until every rank is done (StopAgreement).
        prefill = config.replay_start_size
        self._stop = None
        if self._world > 1:
            prefill = max(prefill, config.minibatch_size)
            self._stop = data_parallel.StopAgreement(window=int(getattr(config, 'dp_stop_window', 64)))
        self._rate_t0, self._rate_steps0 = time.perf_counter(), 0
        self.updates_per_sec = 0.0

        self._prefill_replay_memory(prefill)                      # dqn_agent.py:47
        self._update_target_network()                             # dqn_agent.py:50 (step 0 always syncs)

    # ------------------------------------------------------------------ training loop (dqn_agent.py:52-106)
    def train(self, num_episodes, max_steps_per_episode, supervisor=None):
        for episode in range(num_episodes):
            reward, steps = self.train_episode(max_steps_per_episode)
            self.stats.log_episode(reward, steps)
            logging.info('Episode = %d, steps = %d, reward = %d, training steps = %d, last-100 mean reward = %.2f'
                         % (episode, steps, reward, self.training_steps, self.stats.last_100_mean_reward()))
            self._log_rate()
            if supervisor and supervisor.should_stop():
                logging.warning('Received signal to stop. Exiting train loop.')
                break
        if self._stop is not None:
            # this rank's loop is over; keep the exchange count equal on every rank until all of them are done
            self._stop.finished = True
            while not self._stop.all_finished:
                self._train_minibatch(self.config.minibatch_size)
            if getattr(self.network, 'peers_connected', False):
                self.network.dp_status()
        self.flush_summaries()

    def _log_rate(self):
        """The reference's throughput signal is the Supervisor's 'global_step/sec' line (main.py:136-143); here the same
        number is logged once per episode (updates since the previous line / wall-clock seconds)."""
        now = time.perf_counter()
        dt, dn = now - self._rate_t0, self.training_steps - self._rate_steps0
        if dt > 0 and dn > 0:
            self.updates_per_sec = dn / dt
            logging.info('global_step/sec: %g' % self.updates_per_sec)
        self._rate_t0, self._rate_steps0 = now, self.training_steps

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
        if self._pass_newest:
            self.replay_memory.add(state, action, reward, next_state, done, newest=self.frame_buffer.frames[-1])
        else:
            self.replay_memory.add(state, action, reward, next_state, done)
        return next_state

    # ------------------------------------------------------------------ the hot path (dqn_agent.py:108-143)
    def _train_minibatch(self, minibatch_size):
        rm = self.replay_memory
        if rm.size() < minibatch_size:
            return
        indices = rm.sample_indices(minibatch_size)               # random.sample on the global generator
        want_loss = self._should_log_summary()
        self._last_slot = None
        loss = self._device_step(minibatch_size, indices, want_loss)
        if want_loss:
            if self.async_summary and self._last_slot is not None:
                # graph replay: the loss lands in the slot's pinned buffer when the slot's event fires
                previous, self._pending = self._pending, (self._last_slot, self.training_steps)
                if previous is not None:
                    self._collect(previous)
            elif self.async_summary:
                self._read_loss_later(loss, self.training_steps)
            else:
                self.last_loss = float(loss.item())
                self.summary_writer.add_summary(self.last_loss, self.training_steps)
        elif self._pending is not None:
            self.flush_summaries()
        self.training_steps += 1
        self._update_target_network()
        if self._stop is not None:
            self._stop.after_update(self.training_steps)

    def _read_loss_later(self, loss, step):
        import torch
        if not hasattr(self, '_loss_pin'):
            self._loss_pin = [torch.empty(1, dtype=torch.float32, pin_memory=True) for _ in range(2)]
            self._loss_evt = [torch.cuda.Event(), torch.cuda.Event()]
            self._loss_turn = 0
        t = self._loss_turn
        self._loss_pin[t].copy_(loss, non_blocking=True)          # device -> pinned host, async
        self._loss_evt[t].record()
        previous = self._pending
        self._pending = (t, step)
        self._loss_turn = t ^ 1
        if previous is not None:
            self._collect(previous)

    def _collect(self, pending):
        t, step = pending
        if isinstance(t, dict):                                   # a graph slot
            t['evt'].synchronize()
            self.last_loss = float(t['loss_np'][0])
        else:
            self._loss_evt[t].synchronize()
            self.last_loss = float(self._loss_pin[t].item())
        if self.summary_writer is not None:
            self.summary_writer.add_summary(self.last_loss, step)

    def flush_summaries(self):
        """Writes the summary of the most recent update (pending because summaries trail by one update)."""
        if self._pending is not None:
            pending, self._pending = self._pending, None
            self._collect(pending)
        if self.summary_writer is not None:
            self.summary_writer.flush()

    def _eager_step(self, k, d_idx, want_loss, loss_out=None):
        # single GPU, peer-memory exchange fused into the step's kernels, or NCCL all-reduce (data_parallel.py)
        return data_parallel.train_step(self.network, self.replay_memory, k, d_idx, want_loss=want_loss, loss_out=loss_out)

    def _device_step(self, k, indices, want_loss):
        """One update for the given logical sample indices (host list).  The first call per (k, want_loss) runs eagerly;
        afterwards the step is replayed as a CUDA graph that (1) moves the frames / transitions staged by
        ``ReplayMemory.add`` since the last update from pinned memory into the ring, (2) copies the indices from pinned
        memory, (3) runs gather + update, (4) copies the loss back to pinned memory.  A replay costs the host:
        ``seal`` (header writes), a 128-byte copy and one graph launch.  One graph per pinned staging buffer (three
        rotate), so neither the staging buffer nor the index buffer of the update in flight is overwritten."""
        import torch
        rm = self.replay_memory
        if not self.use_cuda_graph:
            return self._eager_step(k, rm.upload_indices(indices), want_loss)
        key = (k, bool(want_loss))
        g = self._graphs.get(key)
        if g is None:
            self._graphs[key] = {}
            return self._eager_step(k, rm.upload_indices(indices), want_loss)
        stage = rm.seal()                                         # staged frames of buffer `stage` ride along with this update
        other = self._stage_evt[(stage + 1) % 3]
        if other is not None:
            other.synchronize()                                   # adds now refill the next buffer: its last reader (two updates back) must be done
        slot = g.get(stage)
        if slot is None:
            torch.cuda.synchronize()
            pin = torch.empty(k, dtype=torch.int32, pin_memory=True)
            d_idx = torch.empty(k, dtype=torch.int32, device=self.network.device)
            graph = torch.cuda.CUDAGraph()
            loss_pin = torch.zeros(1, dtype=torch.float32, pin_memory=True) if want_loss else None
            with torch.cuda.graph(graph):
                # no copy-engine nodes (several microseconds each inside a graph): the ingest kernel reads the pinned
                # staging buffer and the pinned indices itself, the loss kernel writes the pinned loss word itself
                rm.ingest_staged(stage, pin, d_idx)
                out = self._eager_step(k, d_idx, want_loss, loss_pin)
            slot = dict(graph=graph, out=out, pin=pin, pin_np=pin.numpy(), d_idx=d_idx, evt=torch.cuda.Event(), used=False,
                        loss_pin=loss_pin, loss_np=loss_pin.numpy() if want_loss else None)
            g[stage] = slot
        elif slot['used']:
            slot['evt'].synchronize()                             # the previous replay of this graph has consumed its indices
        slot['pin_np'][:] = indices
        slot['graph'].replay()
        slot['evt'].record()
        slot['used'] = True
        self._stage_evt[stage] = slot['evt']
        self._last_slot = slot
        return slot['out']

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
