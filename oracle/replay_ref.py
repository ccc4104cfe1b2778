"""Pure-Python restatement of the reference's data half (TEST INFRASTRUCTURE).

PINNED against the reference itself: ``oracle/gen_golden.py`` runs the reference's own
``replay_memory.py`` / ``frame_buffer.py`` / ``utils.py`` and stores the results under
``tests/golden/``; ``tests/test_oracle_golden.py`` replays the same scripts through this
file and demands identical output.

Restates:
* ``replay_memory.py:18-53``  ReplayMemory (deque(maxlen) of 5-tuples, random.sample, partition)
* ``frame_buffer.py:13-57``   FrameBuffer (dup-fill on first frame, HWC stack oldest->newest)
* ``utils.py:9-37``           partition, decay_per_step, one_hot
"""
from __future__ import annotations

import collections
import random

import numpy as np


def partition_ref(pred, items):
    """utils.py:9-17 -> (items where pred is false, items where pred is true), both in input order."""
    items = list(items)
    return iter([x for x in items if not pred(x)]), iter([x for x in items if pred(x)])


def one_hot_ref(i, n):
    """utils.py:28-37: float64 vector of length n with a 1 at i."""
    assert i < n, "Invalid args to one_hot"
    v = np.zeros(n)
    v[i] = 1
    return v


def decay_per_step_ref(init_val, min_val, steps):
    """utils.py:22-25."""
    return 0.0 if steps <= 0 else (init_val - min_val) / steps


class ReplayRef:
    """replay_memory.py:18-53.  Logical index 0 is the oldest live transition."""

    def __init__(self, capacity):
        self._cap = capacity
        self.memory = collections.deque(maxlen=capacity)

    def add(self, state, action, reward, observation, terminal):
        self.memory.append((state, action, reward, observation, terminal))

    def size(self):
        return len(self.memory)

    def capacity(self):
        return self._cap

    def sample_indices(self, k):
        """The logical indices ``random.sample(self.memory, k)`` would pick:
        ``random.sample`` only looks at ``len(population)`` and then indexes it."""
        return random.sample(range(len(self.memory)), k)

    def get_minibatch(self, k):
        if len(self.memory) < k:
            return None, None
        picked = [self.memory[i] for i in self.sample_indices(k)]
        return partition_ref(lambda tr: tr[4], picked)

    @staticmethod
    def get_states(it):
        return (tr[0] for tr in it)

    @staticmethod
    def get_next_states(it):
        return (tr[3] for tr in it)


class FrameBufferRef:
    """frame_buffer.py:13-57."""

    def __init__(self, frames_per_state, preprocessor=None):
        if frames_per_state <= 0:
            raise RuntimeError("Frames per state should be greater than 0")
        self.n = frames_per_state
        self.pre = preprocessor or (lambda f: f)
        self.frames = []

    def append(self, frame):
        f = self.pre(frame)
        if not self.frames:
            self.frames = [f] * self.n
        else:
            self.frames = self.frames[1:] + [f]

    def get_state(self):
        if not self.frames:
            return None
        if self.n == 1:
            return self.frames[0]
        return np.stack(self.frames, axis=-1)

    def clear(self):
        self.frames = []
