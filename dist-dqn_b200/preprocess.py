"""Device-side frame preprocessing, batched actor ingest and batched epsilon-greedy acting.

``FramePreprocessor`` is ``utils.resize_image`` (reference src/utils.py:39-45) for batches of frames that are
already on the GPU -- bit-exact with OpenCV, so frames stored through it equal what the reference would have
stored.  ``ingest`` appends the preprocessed frames of many environments to a ReplayMemory's ring without a host
round trip (the reference's FrameBuffer.append + ReplayMemory.add per env step, dqn_agent.py:96-99).
"""
from __future__ import annotations

import numpy as np

import _cabi as _c


class FramePreprocessor(object):
    def __init__(self, src_height, src_width, width, height, device=None):
        import torch
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.src = (int(src_height), int(src_width))
        self.dst = (int(height), int(width))
        h = _c.c_vp()
        _c.check(_c.lib.dqn_preproc_create(_c.C.byref(h), self.src[0], self.src[1], self.dst[0], self.dst[1], self.device),
                 'FramePreprocessor')
        self._h = h

    def __del__(self):
        try:
            if self._h is not None:
                _c.lib.dqn_preproc_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def __call__(self, rgb, out=None):
        """rgb: CUDA uint8 tensor [n, H, W, 3] -> CUDA uint8 tensor [n, height, width]."""
        import torch
        if rgb.dtype != torch.uint8 or rgb.dim() != 4 or tuple(rgb.shape[1:]) != self.src + (3,):
            raise RuntimeError('expected uint8 [n, %d, %d, 3], got %s %s' % (self.src + (tuple(rgb.shape), rgb.dtype)))
        rgb = rgb.contiguous()
        n = rgb.shape[0]
        if out is None:
            out = torch.empty((n,) + self.dst, dtype=torch.uint8, device=rgb.device)
        _c.check(_c.lib.dqn_preproc_run(self._h, rgb.data_ptr(), n, out.data_ptr(), _c.current_stream()), 'preprocess')
        return out


def ingest(replay_memory, frames, actions, rewards, terminals, first_is_episode_start=False):
    """Appends ``n`` consecutive preprocessed frames of ONE environment (CUDA uint8 [n, H, W]) to the ring and
    records the n transitions they complete: transition j goes from the stack ending at the frame before
    ``frames[j]`` to the stack ending at ``frames[j]``.  The episode's first stack must already be in the ring
    (use ``ReplayMemory.add`` for the first step of an episode, which stores its frames_per_state frames)."""
    if replay_memory.handle is None:
        raise RuntimeError('ingest: add the first transition of the episode through ReplayMemory.add first')
    n = int(frames.shape[0])
    pos = _c.c_i64()
    frames = frames.contiguous()
    _c.check(_c.lib.dqn_replay_append_frames_device(replay_memory.handle, frames.data_ptr(), n, _c.C.byref(pos),
                                                    _c.current_stream()), 'ingest')
    p = pos.value
    for j in range(n):
        _c.check(_c.lib.dqn_replay_add(replay_memory.handle, p + j - 1, p + j, int(actions[j]), float(rewards[j]),
                                       1 if terminals[j] else 0), 'ingest')
    replay_memory._resync()                                                # counters moved; the host-side chain is broken
    return p


class BatchedActor(object):
    """E environments stepped in lock-step: the reference's per-step actor work (dqn_agent.py:84-104: pick an action,
    env.step, FrameBuffer.append, ReplayMemory.add) for all of them at once.

    * acting: ONE batched forward + row arg-max over the E current windows (``ReplayMemory.current_states`` ->
      ``Network.predict``) and the reference's epsilon-greedy rule per environment, in environment order: epsilon decays
      BEFORE the roll (dqn_agent.py:162-174), the roll is ``random.random() < epsilon`` on Python's global generator (one
      decay + one roll per environment per step, i.e. the stream E sequential ``_pick_action`` calls would consume), a
      random action is ``env.action_space.sample()``.
    * recording: the E new frames go to the device once (or are already there) and ``ReplayMemory.add_batch`` records the E
      transitions; an environment that finished starts its next episode with ``env.reset()`` on the following step, like
      ``train_episode`` does.
    """

    def __init__(self, envs, network, replay_memory, config, preprocessor=None):
        import torch
        self.envs, self.network, self.rm, self.config = list(envs), network, replay_memory, config
        self.E = len(self.envs)
        self.pre = preprocessor or (lambda f: f)
        self.random_action_prob = config.init_random_action_prob
        self.decay = (config.init_random_action_prob - config.min_random_action_prob) / config.random_action_explore_steps
        self._needs_reset = [True] * self.E
        self._has_window = [False] * self.E
        self._pin = None
        self._dev = None
        self.torch = torch
        self.steps = 0

    def _upload(self, frames):
        torch = self.torch
        arr = np.stack([np.asarray(self.pre(f)) for f in frames])
        if self._pin is None:
            self._pin = torch.empty(arr.shape, dtype=torch.uint8, pin_memory=True)
            self._dev = torch.empty(arr.shape, dtype=torch.uint8, device=self.network.device)
        self._pin.numpy()[...] = arr
        self._dev.copy_(self._pin, non_blocking=True)
        return self._dev

    def pick_actions(self, greedy, stepping):
        """The epsilon-greedy rule of dqn_agent.py:145-174 for the environments that step on this call (``stepping[e]``),
        in environment order; ``greedy``: the E arg-max actions."""
        import random
        acts = [0] * self.E
        for e in range(self.E):
            if not stepping[e]:
                continue
            if self.random_action_prob > self.config.min_random_action_prob:
                self.random_action_prob -= self.decay
            if random.random() < self.random_action_prob:
                acts[e] = int(self.envs[e].action_space.sample())
            else:
                acts[e] = int(greedy[e])
        return acts

    def step(self):
        """One lock-step environment step; returns (rewards, dones) of the environments that stepped (None for those that
        were reset on this call)."""
        E = self.E
        stepping = [not x for x in self._needs_reset]      # the others contribute env.reset()'s observation on this call
        acts = [0] * E
        if any(stepping):
            # every stepping environment has a current window (a freshly reset one got its window on the previous call)
            states = self.rm.current_states() if all(x for x in self._has_window) else None
            _, am = self.network.predict(states, argmax=True)
            acts = self.pick_actions(am.cpu().numpy(), stepping)
        frames, a, r, t, s = [], np.zeros(E, np.int32), np.zeros(E, np.float64), np.zeros(E, np.uint8), np.zeros(E, np.uint8)
        rewards, dones = [None] * E, [None] * E
        for e, env in enumerate(self.envs):
            if self._needs_reset[e]:
                frames.append(env.reset())
                s[e] = 1
                self._needs_reset[e] = False
                self._has_window[e] = True
            else:
                obs, rew, done, _ = env.step(acts[e])
                frames.append(obs)
                a[e], r[e], t[e] = acts[e], rew, 1 if done else 0
                rewards[e], dones[e] = rew, done
                if done:
                    self._needs_reset[e] = True
        self.rm.add_batch(self._upload(frames), a, r, t, s)
        self.steps += 1
        return rewards, dones
