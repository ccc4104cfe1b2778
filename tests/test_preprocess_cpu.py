"""CPU: the preprocessing oracle is pinned against OpenCV itself and the fixture recorded from the reference."""
import os

import numpy as np

from oracle.preprocess_ref import gray_ref, resize_image_ref


def test_restatement_is_bit_exact_with_opencv():
    import cv2
    rng = np.random.default_rng(0)
    for t in range(12):
        shape = (210, 160, 3) if t % 3 else (int(rng.integers(90, 300)), int(rng.integers(90, 300)), 3)
        img = rng.integers(0, 256, shape, dtype=np.uint8)
        assert np.array_equal(gray_ref(img), cv2.cvtColor(img, cv2.COLOR_RGB2GRAY))
        for (w, h) in ((84, 84), (40, 30)):
            assert np.array_equal(resize_image_ref(img, w, h), cv2.resize(cv2.cvtColor(img, cv2.COLOR_RGB2GRAY), (w, h)))


def test_matches_reference_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "utils.npz"))
    assert np.array_equal(resize_image_ref(g["resize_in"], 84, 84), g["resize_out_84"])
    assert list(resize_image_ref(g["resize_in"], 40, 30).shape) == list(g["resize_out_w40_h30_shape"])
