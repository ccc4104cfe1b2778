"""Device-side frame preprocessing and batched actor ingest.

``FramePreprocessor`` is ``utils.resize_image`` (reference src/utils.py:39-45) for batches of frames that are
already on the GPU -- bit-exact with OpenCV, so frames stored through it equal what the reference would have
stored.  ``ingest`` appends the preprocessed frames of many environments to a ReplayMemory's ring without a host
round trip (the reference's FrameBuffer.append + ReplayMemory.add per env step, dqn_agent.py:96-99).
"""
from __future__ import annotations

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
