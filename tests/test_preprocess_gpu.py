"""GPU: device preprocessing is bit-exact with the reference's utils.resize_image (OpenCV), and frames ingested
from device memory come back from the ring exactly like frames added through the host path."""
import os
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_preprocess_bit_exact_vs_reference_fixture_and_oracle(golden_dir):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import preprocess
    from oracle.preprocess_ref import resize_image_ref
    g = np.load(os.path.join(golden_dir, "utils.npz"))
    rng = np.random.default_rng(3)
    imgs = np.concatenate([g["resize_in"][None], rng.integers(0, 256, (7, 210, 160, 3), dtype=np.uint8)])
    pp = preprocess.FramePreprocessor(210, 160, 84, 84)
    out = pp(torch.as_tensor(imgs, device="cuda")).cpu().numpy()
    assert np.array_equal(out[0], g["resize_out_84"])                      # recorded from the reference's own call
    for i in range(len(imgs)):
        assert np.array_equal(out[i], resize_image_ref(imgs[i], 84, 84))
    pp2 = preprocess.FramePreprocessor(210, 160, 40, 30)                    # width 40, height 30 -> [30, 40]
    o2 = pp2(torch.as_tensor(imgs[:2], device="cuda")).cpu().numpy()
    assert o2.shape == (2, 30, 40) and np.array_equal(o2[1], resize_image_ref(imgs[1], 40, 30))
    try:
        import cv2
        assert np.array_equal(out[3], cv2.resize(cv2.cvtColor(imgs[3], cv2.COLOR_RGB2GRAY), (84, 84)))
    except ImportError:
        pass


def test_device_ingest_equals_host_adds():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import preprocess
    import replay_memory
    from frame_buffer import FrameBuffer
    rng = np.random.default_rng(1)
    raw = rng.integers(0, 256, (30, 210, 160, 3), dtype=np.uint8)
    pp = preprocess.FramePreprocessor(210, 160, 84, 84)
    frames_dev = pp(torch.as_tensor(raw, device="cuda"))
    frames = frames_dev.cpu().numpy()
    acts = rng.integers(0, 6, 30); rews = rng.standard_normal(30); terms = np.zeros(30, bool); terms[-1] = True
    host, dev = replay_memory.ReplayMemory(100), replay_memory.ReplayMemory(100)
    fb = FrameBuffer(4)
    fb.append(frames[0]); state = fb.get_state()
    for j in range(1, 30):
        fb.append(frames[j]); nxt = fb.get_state()
        host.add(state, int(acts[j]), float(rews[j]), nxt, bool(terms[j]))
        if j == 1:
            dev.add(state, int(acts[j]), float(rews[j]), nxt, bool(terms[j]))   # first step stores the 4-frame stack
        state = nxt
    preprocess.ingest(dev, frames_dev[2:], acts[2:], rews[2:], terms[2:])
    assert dev.size() == host.size() == 29
    random.seed(4); a = host.sample(16)
    random.seed(4); b = dev.sample(16)
    for x, y in zip(a[:5], b[:5]):
        assert torch.equal(x, y)
