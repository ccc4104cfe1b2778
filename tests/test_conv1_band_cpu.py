"""CPU: the shared-memory budget of the resident conv1 kernels rests on one bound -- a 128-row tile (32 pool windows in
the forward pass's pool-window row order) never needs more than ``FwdResU8::PATCH_ROWS`` = 40 input rows of 336 bytes
(``conv1_band`` in dist-dqn_b200/csrc/umma_core.cuh; the kernels trap if the bound is ever exceeded).  This restates
the device formula and checks the bound for every tile of every batch size up to 70 and a few large ones, including tiles
that straddle two samples (geometry: reference conv1, network.py:318-333 -- 84x84x4 input, 8x8 stride 4 SAME, 2x2 pool)."""
import re
import os

IH, ST, PADH, KS = 84, 4, 2, 8
PH = PW = 11
WPS = PH * PW
TM = 128


def band(m0, B):
    w_first = m0 >> 2
    w_last = min((m0 + TM - 1) >> 2, B * WPS - 1)
    bf, bl = w_first // WPS, w_last // WPS
    pyf, pyl = (w_first - bf * WPS) // PW, (w_last - bl * WPS) // PW
    iy_lo = max(0, 2 * pyf * ST - PADH)
    iy_hi = min(IH - 1, (2 * pyl + 1) * ST - PADH + KS - 1)
    return bf * IH + iy_lo, bl * IH + iy_hi


def _patch_rows():
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dist-dqn_b200", "csrc",
                            "umma_core.cuh")).read()
    return int(re.search(r"static constexpr int PATCH_ROWS = (\d+);", src).group(1))


def test_band_never_exceeds_the_patch():
    limit = _patch_rows()
    worst = 0
    for B in list(range(1, 71)) + [128, 130, 512, 513]:
        for m0 in range(0, B * WPS * 4, TM):
            lo, hi = band(m0, B)
            assert lo <= hi
            worst = max(worst, hi - lo + 1)
    assert worst <= limit, (worst, limit)
    assert worst >= 30          # the formula is exercised (5 pool rows -> 8*4+4 = 36 input rows)
