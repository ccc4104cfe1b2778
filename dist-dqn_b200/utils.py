"""Helpers with the reference's names (reference src/utils.py:9-45).  ``partition`` and ``one_hot``
exist for callers that still use the tuple-returning ``ReplayMemory.get_minibatch``; the device
step never builds one-hot vectors (the action index goes straight into the fused TD kernel)."""
import numpy as np


def partition(pred, iterable):
    """(items failing pred, items passing pred), each lazily, in input order (utils.py:9-17)."""
    pool = list(iterable)
    return (x for x in pool if not pred(x)), (x for x in pool if pred(x))


def decay(val, min_val, decay_rate):
    return max(val * decay_rate, min_val)


def decay_per_step(init_val, min_val, steps):
    """Per-step decrement of a linear schedule; 0.0 when steps <= 0 (utils.py:22-25)."""
    if steps <= 0:
        return 0.0
    return (init_val - min_val) / float(steps)


def one_hot(i, n):
    """float64 vector of length n, 1 at position i (utils.py:28-37)."""
    assert i < n, "Invalid args to one_hot"
    vec = np.zeros(n)
    vec[i] = 1
    return vec


def resize_image(image, width, height):
    """RGB -> gray -> (height, width) uint8, bilinear (utils.py:39-45).  Actor-side; uses OpenCV
    like the reference."""
    import cv2
    return cv2.resize(cv2.cvtColor(image, cv2.COLOR_RGB2GRAY), (width, height))
