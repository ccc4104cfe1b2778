"""FrameBuffer with the reference's surface (reference src/frame_buffer.py:13-57).

Actor-side helper: keeps the last ``frames_per_state`` preprocessed frames and hands out the
current state as an HWC stack whose channel axis is time, oldest first.  The training side never
stores these stacks: ``ReplayMemory.add`` recognises consecutive stacks and keeps one copy of each
frame in the device ring (see replay_memory.py), so this class stays a small host object.
"""
import numpy as np

try:                      # OpenCV is already a dependency of the reference (utils.resize_image); its channel merge is
    import cv2            # a much faster interleave than np.stack for uint8 frames
except ImportError:       # pragma: no cover
    cv2 = None


def _identity(frame):
    return frame


def _stack_last(frames):
    """np.stack(frames, axis=-1), through cv2.merge when that yields the identical array."""
    f0 = frames[0]
    if (cv2 is not None and isinstance(f0, np.ndarray) and f0.ndim == 2 and f0.dtype == np.uint8
            and 2 <= len(frames) <= 4):
        return cv2.merge(frames)
    return np.stack(frames, axis=-1)


class FrameBuffer(object):
    def __init__(self, frames_per_state, preprocessor=_identity):
        if frames_per_state <= 0:
            raise RuntimeError('Frames per state should be greater than 0')
        self.frames_per_state = frames_per_state
        self.preprocessor = preprocessor
        self._window = []          # oldest ... newest, either empty or exactly frames_per_state long

    @property
    def frames(self):
        """The live window (the reference exposes its deque under this name)."""
        return self._window

    def append(self, frame):
        """Push one raw frame; the very first frame after clear() fills the whole window
        (frame_buffer.py:35-36), later frames shift it by one."""
        processed = self.preprocessor(frame)
        if self._window:
            del self._window[0]
            self._window.append(processed)
        else:
            self._window = [processed for _ in range(self.frames_per_state)]

    def get_state(self):
        """None when empty, the bare frame when frames_per_state == 1, else a fresh
        ``[..., frames_per_state]`` array (frame_buffer.py:47-51)."""
        if not self._window:
            return None
        if self.frames_per_state == 1:
            return self._window[0]
        return _stack_last(self._window)

    def clear(self):
        self._window = []
