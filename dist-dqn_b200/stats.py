"""Episode statistics with the reference's surface (reference src/stats.py:7-19). Host-only, not
on the hot path; one instance per agent (not thread-safe, like the reference's)."""
import collections


class Stats(object):
    WINDOW = 100

    def __init__(self):
        self.rewards = collections.deque(maxlen=self.WINDOW)
        self.episodes = 0
        self.total_steps = 0

    def log_episode(self, reward, steps):
        self.rewards.append(reward)
        self.episodes += 1
        self.total_steps += steps

    def last_100_mean_reward(self):
        n = len(self.rewards)
        return float('nan') if n == 0 else float(sum(self.rewards)) / n
