"""CPU: the MMA schedules of the space-to-depth conv1 kernels (dist-dqn_b200/csrc/conv1_s2d.cuh), read from the source.

conv1 as a 2x2 convolution over super-pixels (network.py:395-407 restated in DESIGN.md section 4): the conv output at
pool-window position (wy, wx) takes filter tap (dy, dx) from the image slice at super-pixel offset (wy+dy, wx+dx).  The
kernels merge MMAs that read the same slice into wider ones; these tests check that every (position, tap) pair is
still computed exactly once, into the right accumulator, with the right weights / gradient tile."""
import os
import re

from conftest import ROOT

SRC = open(os.path.join(ROOT, "dist-dqn_b200", "csrc", "conv1_s2d.cuh")).read()
FWD = SRC[SRC.index("conv1_s2d_fwd_kernel(FwdArgs a)"):SRC.index("// ---------------------------------------------------------------- epilogue: TMEM lane quarter")]
WG = SRC[SRC.index("conv1_s2d_wgrad_kernel(WgradArgs a)"):SRC.index("// ---------------------------------------------------------------- builders: image + routed gradient tiles")]


def test_forward_schedule_covers_every_window_position_and_tap_once():
    order = [int(v) for v in re.search(r"order\[FWD_W_BLOCKS\] = \{([0-9, ]+)\}", FWD).group(1).split(",")]
    assert len(order) == 8 and sorted(order[:4]) == sorted(order[4:]) == [0, 1, 2, 3]      # block b holds tap order[b] = 2*dy + dx
    runs = re.findall(r"run\((\d), (\d), (\d), (\d+), idesc(\d+), (true|false)\);", FWD)
    assert len(runs) == 10
    window_of_colblock = [0, 1, 3, 2]                 # accumulator column order [w00, w01, w11, w10], w = 2*wy + wx
    # the epilogue reads window position w from column block w ^ (w >> 1): the same order
    assert "(w ^ (w >> 1)) * 32" in SRC and [w ^ (w >> 1) for w in range(4)] == [window_of_colblock.index(w) for w in range(4)]
    seen = {}
    mmas = 0
    for k, (Y, X, wblock, col, n, first) in enumerate(runs):
        Y, X, wblock, col, n = int(Y), int(X), int(wblock), int(col), int(n)
        assert col % 32 == 0 and n in (32, 64, 128) and col + n <= 128 and wblock + n // 32 <= 8
        assert (first == "true") == (k == 0)
        if k == 0:
            assert n == 128 and col == 0              # the first run initialises all four accumulators
        mmas += 4                                     # 64 channels = 4 k-steps of 16
        for i in range(n // 32):
            w, tap = window_of_colblock[col // 32 + i], order[wblock + i]
            wy, wx, dy, dx = w >> 1, w & 1, tap >> 1, tap & 1
            assert (wy + dy, wx + dx) == (Y, X), (k, i)
            assert (w, tap) not in seen
            seen[(w, tap)] = k
    assert len(seen) == 16 and mmas == 40


def test_weight_gradient_schedule_covers_every_window_position_and_tap_column_once():
    runs = re.findall(r"run\((\d), (\d), (\d), idesc(\d+), (true|false)\);", WG)
    assert len(runs) == 3                             # per wy; the kernel loops wy = 0, 1 around them
    assert "tmem + wy * 64 + col * 32" in WG and "(wy * 2 + wx) * 8192" in WG
    # epilogue: accumulator (wy, dx) sits in column block wy*2 + (1 - dx)
    assert "(2 + (1 - dxe)) * 32" in SRC and "(1 - dxe) * 32, acc0" in SRC
    seen = set()
    for k, (t, wx, col, n, first) in enumerate(runs):
        t, wx, col, n = int(t), int(wx), int(col), int(n)
        assert (first == "true") == (k == 0)
        if k == 0:
            assert n == 64 and col == 0               # the first run initialises both accumulators of this wy
        for i in range(n // 32):
            wxi, dx = wx + i, 1 - (col + i)           # B tile dZ(wy, wx+i) is the neighbour in shared memory
            assert 0 <= wxi <= 1 and 0 <= dx <= 1 and wxi + dx == t, (k, i)
            assert (wxi, dx) not in seen
            seen.add((wxi, dx))
    assert len(seen) == 4


def test_weight_image_permutation_round_trips():
    """kperm (reference reduction index k = (ky*8+kx)*4+f -> image index) and kref (tap, channel -> k) are inverse."""
    def kperm(k):
        return (((k >> 5) >> 2) * 2 + (((k >> 2) & 7) >> 2)) * 64 + ((k >> 5) & 3) * 16 + ((k >> 2) & 3) * 4 + (k & 3)

    def kref(dy, dx, c64):
        return ((4 * dy + (c64 >> 4)) * 8 + 4 * dx + ((c64 >> 2) & 3)) * 4 + (c64 & 3)
    assert re.search(r"return \(\(\(k >> 5\) >> 2\) \* 2 \+ \(\(\(k >> 2\) & 7\) >> 2\)\) \* 64 \+ \(\(k >> 5\) & 3\) \* 16 \+ \(\(k >> 2\) & 3\) \* 4 \+ \(k & 3\);", SRC)
    assert re.search(r"return \(\(4 \* dy \+ \(c64 >> 4\)\) \* 8 \+ 4 \* dx \+ \(\(c64 >> 2\) & 3\)\) \* 4 \+ \(c64 & 3\);", SRC)
    assert sorted(kperm(k) for k in range(256)) == list(range(256))
    for dy in range(2):
        for dx in range(2):
            for c in range(64):
                k = kref(dy, dx, c)
                assert kperm(k) == (2 * dy + dx) * 64 + c
                ky, kx, f = k >> 5, (k >> 2) & 7, k & 3           # pixel (4*dy + yy, 4*dx + xx) of frame f
                assert (ky >> 2, kx >> 2) == (dy, dx) and ((ky & 3) * 4 + (kx & 3)) * 4 + f == c
