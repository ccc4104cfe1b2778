"""ReplayMemory with the reference's surface (reference src/replay_memory.py:18-53) on top of the
device-resident frame ring of libdqn_b200 (csrc/replay.cu).

What changes underneath
* The reference keeps (state, action, reward, next_state, terminal) tuples in a deque, i.e. every
  frame four times in host RAM, and samples with ``random.sample(deque, k)`` (O(n) indexing).
* Here ``add`` recognises that ``state`` is the previous transition's ``next_state`` and that
  ``next_state`` is ``state`` shifted by one frame, and appends only the NEW frame to a ring in HBM
  (one copy per frame).  Stacks that do not chain (episode start, arbitrary callers) are simply
  stored in full, so any input sequence round-trips bit-exactly.
* ``get_minibatch`` draws ``random.sample(range(size), k)`` from Python's global generator -- the
  same logical indices ``random.sample(self.memory, k)`` picks for the same seed -- and one CUDA
  gather kernel rebuilds the stacks.  ``sample`` is the device-resident fast path the agent uses.
"""
from __future__ import annotations

import collections
import random

import numpy as np

import _cabi as _c

Batch = collections.namedtuple("Batch", "states next_states actions rewards terminals indices")
_PUSH = _c.lib.dqn_replay_push


class ReplayMemory(object):
    def __init__(self, capacity, device=None, frame_capacity=0, trusted=False):
        """capacity as in the reference; the rest are optional extensions: CUDA device index
        (default: current), frame ring size (default 2*capacity+2*frames_per_state), and
        ``trusted`` to skip the content comparison in ``add`` when ``state is previous next_state``."""
        self._capacity = int(capacity)
        self._device = device
        self._frame_capacity = int(frame_capacity)
        self._trusted = trusted
        self._h = None
        self._last_next = None       # the ndarray passed as next_state on the previous add
        self._last_next_pos = -1
        self._stack = None           # (frames_per_state, frame_shape, dtype) once known
        self._bufs = {}
        self._size = 0               # host mirror of the ring's counters (saves ctypes round trips on the hot path)
        self._frames_total = 0
        self._scratch = None
        self._stage_out = _c.c_i32()
        self._stage_ref = _c.C.byref(self._stage_out)

    # ------------------------------------------------------------------ construction of the ring
    @staticmethod
    def _describe(state):
        a = np.asarray(state)
        if a.ndim == 3 and a.shape[-1] <= 16:
            return a.shape[-1], a.shape[:-1], a.dtype          # HWC stack, channel = time
        return 1, a.shape, a.dtype                             # bare frame / vector observation

    def _ensure(self, state):
        if self._h is not None:
            return
        import torch
        S, fshape, dt = self._describe(state)
        if dt.itemsize not in (1, 4, 8):
            raise RuntimeError("unsupported observation dtype %s" % dt)
        dev = torch.cuda.current_device() if self._device is None else int(self._device)
        self._device = dev
        self._stack = (S, tuple(fshape), np.dtype(dt))
        self._elems = int(np.prod(fshape)) if len(fshape) else 1
        h = _c.c_vp()
        _c.check(_c.lib.dqn_replay_create(_c.C.byref(h), self._capacity, self._frame_capacity,
                                          self._elems * dt.itemsize, dt.itemsize, S, dev), "ReplayMemory")
        self._h = h

    def __del__(self):
        try:
            if self._h is not None:
                _c.lib.dqn_replay_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    @property
    def device(self):
        return self._device

    # ------------------------------------------------------------------ reference surface
    def _frames_of(self, stack):
        """[S, frame...] contiguous view/copy of an HWC stack (or the bare frame as [1, ...])."""
        S, fshape, dt = self._stack
        a = np.asarray(stack)
        if a.dtype != dt or tuple(a.shape) != (fshape + (S,) if S > 1 else fshape):
            raise RuntimeError("observation shape/dtype changed: got %s %s" % (a.shape, a.dtype))
        if S == 1:
            return np.ascontiguousarray(a).reshape((1,) + fshape)
        return np.ascontiguousarray(np.moveaxis(a, -1, 0))

    def _append(self, frames):
        pos = _c.c_i64()
        _c.check(_c.lib.dqn_replay_append_frames(self._h, frames.ctypes.data, frames.shape[0], _c.C.byref(pos)),
                 "ReplayMemory.add")
        self._frames_total = pos.value + frames.shape[0]
        return pos.value

    def add(self, state, action, reward, observation, terminal, newest=None):
        """replay_memory.py:22-23.  ``observation`` is the next state.  ``newest`` (optional extension, trusted mode
        only): the newest frame of ``observation`` as its own contiguous array -- what ``FrameBuffer.frames[-1]`` holds --
        which saves the strided extraction of ``observation[..., -1]``."""
        if (self._trusted and state is self._last_next and self._h is not None
                and self._last_next_pos == self._frames_total - 1):
            # fast lane (the agent loop, dqn_agent.py:96-100): one new frame, one C call
            if self._scratch is not None:
                np.copyto(self._scratch, observation[..., -1] if newest is None else newest)
                ptr = self._scratch_ptr                      # cached: ndarray.ctypes is a slow attribute
            else:
                ptr = np.ascontiguousarray(observation).ctypes.data
            rc = _PUSH(self._h, ptr, int(action), float(reward), 1 if terminal else 0)
            if rc:
                _c.check(rc, "ReplayMemory.add")
            self._last_next = observation
            self._last_next_pos += 1
            self._frames_total += 1
            if self._size < self._capacity:
                self._size += 1
            return
        self._ensure(state)
        S = self._stack[0]
        if self._trusted:
            chained = state is self._last_next
        else:   # _last_next is a private copy, so in-place edits by the caller cannot fool the check
            chained = self._last_next is not None and np.array_equal(state, self._last_next)
        if chained:
            spos = self._last_next_pos
        else:
            fr = self._frames_of(state)
            spos = self._append(fr) + S - 1
        nxt = np.asarray(observation)
        nfp = _c.c_i64()
        _c.lib.dqn_replay_next_frame_pos(self._h, _c.C.byref(nfp))
        shifted = (nfp.value == spos + 1) and (
            S == 1 or (self._trusted and chained) or np.array_equal(nxt[..., :-1], np.asarray(state)[..., 1:]))
        if shifted:
            newest = nxt[..., -1] if S > 1 else nxt
            npos = self._append(np.ascontiguousarray(newest).reshape((1,) + self._stack[1]))
        else:
            npos = self._append(self._frames_of(nxt)) + S - 1
        _c.check(_c.lib.dqn_replay_add(self._h, spos, npos, int(action), float(reward), 1 if terminal else 0),
                 "ReplayMemory.add")
        self._last_next = observation if self._trusted else np.array(observation, copy=True)
        self._last_next_pos = npos
        if self._size < self._capacity:
            self._size += 1
        if self._scratch is None and S > 1:
            self._scratch = np.empty(self._stack[1], self._stack[2])
            self._scratch_ptr = self._scratch.ctypes.data

    def size(self):
        return self._size

    def device_size(self):
        """The ring's own counter (equals size(); used by tests and after device-side ingest)."""
        if self._h is None:
            return 0
        n = _c.c_i64()
        _c.lib.dqn_replay_size(self._h, _c.C.byref(n), None)
        return n.value

    def _resync(self):
        """Re-reads the counters from the library after frames / transitions were added behind the shell's back."""
        n, p = _c.c_i64(), _c.c_i64()
        _c.lib.dqn_replay_size(self._h, _c.C.byref(n), None)
        _c.lib.dqn_replay_next_frame_pos(self._h, _c.C.byref(p))
        self._size, self._frames_total = n.value, p.value
        self._last_next, self._last_next_pos = None, -1

    def capacity(self):
        return self._capacity

    def get_minibatch(self, minibatch_size):
        """replay_memory.py:31-45: (non_terminal_iter, terminal_iter) of 5-tuples in sample order,
        or (None, None) when fewer than ``minibatch_size`` transitions are stored.  Compatibility
        path: copies the gathered batch back to the host."""
        if self.size() < minibatch_size:
            return None, None
        b = self.sample(minibatch_size)
        states = b.states.cpu().numpy()
        nexts = b.next_states.cpu().numpy()
        acts = b.actions.cpu().numpy()
        rews = b.rewards.cpu().numpy()
        terms = b.terminals.cpu().numpy()
        items = [(states[i], int(acts[i]), float(rews[i]), nexts[i], bool(terms[i])) for i in range(minibatch_size)]
        return (x for x in items if not x[4]), (x for x in items if x[4])

    @staticmethod
    def get_states(iterable):
        return (item[0] for item in iterable)

    @staticmethod
    def get_next_states(iterable):
        return (item[3] for item in iterable)

    # ------------------------------------------------------------------ device-resident fast path
    def sample_indices(self, k):
        """The logical indices (0 = oldest) ``random.sample(self.memory, k)`` would pick in the
        reference for the current state of Python's global generator."""
        return random.sample(range(self.size()), k)

    def _buffers(self, k):
        import torch
        if k not in self._bufs:
            S, fshape, dt = self._stack
            tdt = {np.dtype(np.uint8): torch.uint8, np.dtype(np.float32): torch.float32,
                   np.dtype(np.float64): torch.float64, np.dtype(np.int32): torch.int32,
                   np.dtype(np.int64): torch.int64}.get(dt)
            if tdt is None:
                raise RuntimeError("unsupported observation dtype %s" % dt)
            dev = torch.device("cuda", self._device)
            shp = (k,) + fshape + ((S,) if S > 1 else ())
            self._bufs[k] = dict(
                idx_pin=[torch.empty(k, dtype=torch.int32, pin_memory=True) for _ in range(4)],
                idx_evt=[None] * 4, idx_turn=0,
                idx=torch.empty(k, dtype=torch.int32, device=dev),
                states=torch.empty(shp, dtype=tdt, device=dev), next_states=torch.empty(shp, dtype=tdt, device=dev),
                actions=torch.empty(k, dtype=torch.int32, device=dev),
                rewards=torch.empty(k, dtype=torch.float64, device=dev),
                terminals=torch.empty(k, dtype=torch.uint8, device=dev))
        return self._bufs[k]

    def upload_indices(self, indices):
        """Host list of logical indices -> device int32 tensor (pinned staging, async)."""
        k = len(indices)
        b = self._buffers(k)
        import torch
        t = b["idx_turn"]
        b["idx_turn"] = (t + 1) % 4
        if b["idx_evt"][t] is not None:
            b["idx_evt"][t].synchronize()      # the pinned buffer may still be the source of an earlier copy
        b["idx_pin"][t].numpy()[:] = indices
        b["idx"].copy_(b["idx_pin"][t], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        b["idx_evt"][t] = ev
        return b["idx"]

    def gather(self, d_indices, out=None):
        """Gathers the transitions at the given device int32 logical indices into device tensors."""
        k = int(d_indices.numel())
        b = out or self._buffers(k)
        _c.check(_c.lib.dqn_replay_gather(self._h, _c.ptr(d_indices), k, _c.ptr(b["states"]), _c.ptr(b["next_states"]),
                                          _c.ptr(b["actions"]), _c.ptr(b["rewards"]), _c.ptr(b["terminals"]),
                                          _c.current_stream()), "ReplayMemory.gather")
        return Batch(b["states"], b["next_states"], b["actions"], b["rewards"], b["terminals"], d_indices)

    def sample(self, k, indices=None):
        """Device-resident minibatch: a ``Batch`` of CUDA tensors (states/next_states keep the
        observation dtype, e.g. uint8 [k,84,84,4])."""
        if self.size() < k:
            raise RuntimeError("ReplayMemory.sample: only %d transitions stored, %d requested" % (self.size(), k))
        if indices is None:
            indices = self.sample_indices(k)
        return self.gather(self.upload_indices(indices))

    # ------------------------------------------------------------------ lock-step actors on the device
    def set_lanes(self, num_envs, frame_shape=(84, 84), frames_per_state=4, dtype=np.uint8):
        """Switches the (empty) memory to lock-step actor mode: ``num_envs`` environments add one frame each per call of
        ``add_batch`` and no frame passes through the host.  Creates the ring if it does not exist yet."""
        if self._h is None:
            probe = np.zeros(tuple(frame_shape) + ((frames_per_state,) if frames_per_state > 1 else ()), dtype)
            self._ensure(probe)
        _c.check(_c.lib.dqn_replay_set_lanes(self._h, int(num_envs)), "ReplayMemory.set_lanes")
        self._lanes = int(num_envs)

    def add_batch(self, frames, actions, rewards, terminals, episode_start):
        """One preprocessed frame per environment (CUDA uint8 tensor [E, H, W], environment order) plus host sequences of E
        actions / rewards / terminals / episode-start flags: what E calls of ``FrameBuffer.append`` + ``ReplayMemory.add``
        do in the reference (dqn_agent.py:96-100), in one call.  A frame flagged ``episode_start`` is ``env.reset()``'s
        observation: it starts a new window (the frame repeated frames_per_state times) and records no transition."""
        E = self._lanes
        frames = frames.contiguous()
        if int(frames.shape[0]) != E:
            raise RuntimeError("add_batch: expected %d frames, got %d" % (E, int(frames.shape[0])))
        a = np.ascontiguousarray(actions, np.int32)
        r = np.ascontiguousarray(rewards, np.float64)
        t = np.ascontiguousarray(terminals, np.uint8)
        s = np.ascontiguousarray(episode_start, np.uint8)
        _c.check(_c.lib.dqn_replay_add_batch(self._h, E, frames.data_ptr(), a.ctypes.data, r.ctypes.data, t.ctypes.data,
                                             s.ctypes.data, _c.current_stream()), "ReplayMemory.add_batch")
        self._size = min(self._capacity, self._size + int(E - int(s.astype(bool).sum())))
        self._last_next, self._last_next_pos = None, -1

    def current_states(self, out=None):
        """The E environments' current windows as a CUDA tensor [E, H, W, frames_per_state] (``FrameBuffer.get_state()``
        of every environment), ready for ``Network.predict``."""
        import torch
        S, fshape, dt = self._stack
        if out is None:
            out = torch.empty((self._lanes,) + fshape + ((S,) if S > 1 else ()), dtype=torch.uint8,
                              device=torch.device("cuda", self._device))
        _c.check(_c.lib.dqn_replay_current_states(self._h, out.data_ptr(), _c.current_stream()), "ReplayMemory.current_states")
        return out

    def flush(self):
        if self._h is not None:
            _c.check(_c.lib.dqn_replay_flush(self._h, _c.current_stream()), "ReplayMemory.flush")

    # graph-embedded ingest (see include/dqn_b200.h): the step graph itself moves the staged frames to the ring
    GRAPH_INGEST_FRAMES = 8

    def seal(self):
        """Host-only: publishes the current staging buffer for the step graph and switches to the other one;
        returns the index of the sealed buffer."""
        rc = _c.lib.dqn_replay_seal(self._h, self._stage_ref, None)     # the stream only matters for the eager fallback
        if rc:
            _c.check(rc, "ReplayMemory.seal")
        return self._stage_out.value

    def ingest_staged(self, stage, pinned_idx=None, d_idx=None):
        """Enqueues (capturable) the kernel that scatters staging buffer ``stage`` into the ring; optionally the same
        kernel copies the sample indices from a pinned host tensor to a device tensor."""
        k = 0 if pinned_idx is None else int(pinned_idx.numel())
        _c.check(_c.lib.dqn_replay_ingest_staged(self._h, int(stage), _c.ptr(pinned_idx), _c.ptr(d_idx), k,
                                                 _c.current_stream()), "ReplayMemory.ingest_staged")


class IndexTable(object):
    """Sample indices of the next `rows` updates, drawn at once on the host (``random.sample`` per update, the
    reference's sampler, dqn_agent.py:110) and kept on the device: ``next(d_idx)`` enqueues ONE kernel that copies
    the current row into ``d_idx`` and advances a device-side cursor, so a captured update replays without the host
    touching the indices again (dqn_index_table_next)."""

    def __init__(self, rows_of_indices, device):
        import torch
        if isinstance(rows_of_indices, torch.Tensor):
            self.table = rows_of_indices.to(device=device, dtype=torch.int32).contiguous()
        else:
            self.table = torch.tensor(rows_of_indices, dtype=torch.int32, device=device).contiguous()
        self.rows, self.k = int(self.table.shape[0]), int(self.table.shape[1])
        self.cursor = torch.zeros(1, dtype=torch.int64, device=device)
        self.scratch = torch.empty(self.k, dtype=torch.int32, device=device)     # for kernels that take the row as a plain array

    def next(self, d_idx):
        _c.check(_c.lib.dqn_index_table_next(_c.ptr(self.table), self.rows, self.k, _c.ptr(self.cursor), _c.ptr(d_idx),
                                             _c.current_stream()), "IndexTable.next")
        return d_idx
