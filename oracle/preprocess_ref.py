"""numpy restatement of the reference's frame preprocessing (TEST INFRASTRUCTURE).

``utils.resize_image`` (reference src/utils.py:39-45) = ``cv2.cvtColor(RGB2GRAY)`` followed by ``cv2.resize`` with the
default INTER_LINEAR.  OpenCV (4.13 here; a third-party dependency the reference does not pin) does both in fixed
point for uint8; the arithmetic below restates OpenCV's published algorithm:
  gray  = (R*9798 + G*19235 + B*3735 + 2^14) >> 15
  resize: source coordinate f = (d + 0.5) * (src/dst) - 0.5, clamped to the image; tap weights rounded to 11 bits
          (cvRound(w * 2048)); horizontal pass in int32, vertical pass
          dst = (((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2.
PINNED: checked bit-exactly against cv2 itself (tests/test_preprocess_cpu.py) and against
``tests/golden/utils.npz`` recorded from the reference's own ``utils.resize_image``.
"""
import numpy as np


def gray_ref(img):
    r, g, b = (img[..., i].astype(np.int64) for i in range(3))
    return ((r * 9798 + g * 19235 + b * 3735 + 16384) >> 15).astype(np.uint8)


def linear_taps(dn, sn):
    """(left source index, weight of left tap, weight of right tap) per destination coordinate, 11-bit weights."""
    idx = np.zeros(dn, np.int64)
    a0 = np.zeros(dn, np.int64)
    a1 = np.zeros(dn, np.int64)
    scale = float(sn) / float(dn)
    for d in range(dn):
        f = (d + 0.5) * scale - 0.5
        s = int(np.floor(f))
        f -= s
        if s < 0:
            s, f = 0, 0.0
        if s >= sn - 1:
            s, f = sn - 1, 0.0
        idx[d] = s
        f32 = np.float32(f)
        a0[d] = int(np.rint((np.float32(1.0) - f32) * np.float32(2048)))
        a1[d] = int(np.rint(f32 * np.float32(2048)))
    return idx, a0, a1


def resize_ref(src, width, height):
    sh, sw = src.shape
    xi, xa0, xa1 = linear_taps(width, sw)
    yi, yb0, yb1 = linear_taps(height, sh)
    s = src.astype(np.int64)
    rows = s[:, xi] * xa0 + s[:, np.minimum(xi + 1, sw - 1)] * xa1
    s0, s1 = rows[yi], rows[np.minimum(yi + 1, sh - 1)]
    out = (((yb0[:, None] * (s0 >> 4)) >> 16) + ((yb1[:, None] * (s1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


def resize_image_ref(image, width, height):
    """utils.resize_image: returns a (height, width) uint8 frame."""
    return resize_ref(gray_ref(image), width, height)
