"""Minimal stand-in for the old-gym API the reference drives (reference src/dqn_agent.py:78, 87, 156;
src/main.py:101, 125): ``reset()``, ``step(a) -> (obs, reward, done, info)``, ``action_space.n /
.sample()``, ``observation_space.shape``, ``spec.id``.  gym itself is not installable on this image
(SURVEY.md F3); if it is importable it is used instead.  CartPole follows the classic cart-pole
constants gym published (env dynamics parity with gym is unpinned); the Atari stand-in emits
synthetic frames because ROMs are not available offline.
"""
import math

import numpy as np


class _Space(object):
    def __init__(self, n=None, shape=None, rng=None):
        self.n, self.shape, self._rng = n, shape, rng

    def sample(self):
        return int(self._rng.integers(0, self.n))


class _Spec(object):
    def __init__(self, env_id):
        self.id = env_id


class _NullMonitor(object):
    def start(self, *a, **k):
        pass

    def close(self):
        pass


class CartPoleEnv(object):
    """Classic cart-pole: gravity 9.8, cart 1.0 kg, pole 0.1 kg, half-length 0.5 m, force 10 N,
    tau 0.02 s Euler, fails at |x|>2.4 or |theta|>12 deg, reward 1 per step, 200-step limit."""

    def __init__(self, env_id='CartPole-v0', seed=None, max_steps=200):
        self.spec = _Spec(env_id)
        self._rng = np.random.default_rng(seed)
        self.action_space = _Space(n=2, rng=self._rng)
        self.observation_space = _Space(shape=(4,))
        self.monitor = _NullMonitor()
        self._max_steps = max_steps
        self._s = None
        self._t = 0

    def reset(self):
        self._s = self._rng.uniform(-0.05, 0.05, 4)
        self._t = 0
        return self._s.copy()

    def step(self, action):
        x, x_dot, th, th_dot = self._s
        force = 10.0 if action == 1 else -10.0
        cos, sin = math.cos(th), math.sin(th)
        total_mass, pml = 1.1, 0.05
        temp = (force + pml * th_dot * th_dot * sin) / total_mass
        th_acc = (9.8 * sin - cos * temp) / (0.5 * (4.0 / 3.0 - 0.1 * cos * cos / total_mass))
        x_acc = temp - pml * th_acc * cos / total_mass
        x, x_dot = x + 0.02 * x_dot, x_dot + 0.02 * x_acc
        th, th_dot = th + 0.02 * th_dot, th_dot + 0.02 * th_acc
        self._s = np.array([x, x_dot, th, th_dot])
        self._t += 1
        done = bool(abs(x) > 2.4 or abs(th) > 12 * 2 * math.pi / 360 or self._t >= self._max_steps)
        return self._s.copy(), 1.0, done, {}


class SyntheticAtariEnv(object):
    """Atari-shaped stand-in: uint8 RGB 210x160 frames (or pre-sized gray frames), random episode
    lengths, rewards in {-1,0,1}.  For throughput runs only -- nothing can be learnt from it."""

    def __init__(self, env_id='SyntheticPong-v0', num_actions=6, seed=None, frame_shape=(210, 160, 3),
                 min_len=50, max_len=2000, pool=64):
        self.spec = _Spec(env_id)
        self._rng = np.random.default_rng(seed)
        self.action_space = _Space(n=num_actions, rng=self._rng)
        self.observation_space = _Space(shape=tuple(frame_shape))
        self.monitor = _NullMonitor()
        self._pool = self._rng.integers(0, 256, (pool,) + tuple(frame_shape), dtype=np.uint8)
        self._min, self._max = min_len, max_len
        self._left = 0

    def _frame(self):
        return self._pool[int(self._rng.integers(0, len(self._pool)))]

    def reset(self):
        self._left = int(self._rng.integers(self._min, self._max + 1))
        return self._frame()

    def step(self, action):
        self._left -= 1
        return self._frame(), float(self._rng.integers(-1, 2)), self._left <= 0, {}


def make(env_id, seed=None):
    try:
        import gym                                   # the reference's path (main.py:101)
        return gym.make(env_id)
    except ImportError:
        pass
    if env_id.startswith('CartPole'):
        return CartPoleEnv(env_id, seed=seed)
    low = env_id.lower()
    if 'breakout' in low:
        return SyntheticAtariEnv(env_id, num_actions=4, seed=seed)
    if 'pong' in low or 'synthetic' in low or 'atari' in low:
        return SyntheticAtariEnv(env_id, num_actions=6, seed=seed)
    raise RuntimeError('unknown environment %s (gym is not installed; built-in: CartPole-*, *Pong*, *Breakout*)' % env_id)
