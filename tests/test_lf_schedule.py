"""Loop-filter dependency schedule (vp9_lf.h: lf_schedule_dir) == the reference's sequential edge
order (vp9lpf.c:31-181), on random and on structured VP9Filter masks.  CPU only."""
import ctypes as C

import numpy as np
import pytest

from common import model_lib, ptr


def _check(lib, filt, chroma, first_col, first_row, seed, bpp):
    steps = C.c_int(0)
    r = lib.model_lf_sched_check(ptr(filt), chroma, first_col, first_row, seed, bpp, C.byref(steps))
    return r, steps.value


@pytest.mark.parametrize("bpp", [8, 10, 12])
def test_schedule_random_masks(bpp):
    lib = model_lib()
    rng = np.random.default_rng(1234 + bpp)
    for it in range(300):
        filt = rng.integers(0, 256, 192, dtype=np.uint8)
        if it % 3 == 0:                      # sparser masks: realistic edge density
            filt[64:] &= rng.integers(0, 256, 128, dtype=np.uint8)
        filt[:64] = rng.integers(0, 64, 64)
        for chroma in (0, 1):
            r, _ = _check(lib, filt, chroma, it & 1, (it >> 1) & 1, it, bpp)
            assert r == 0, (it, chroma, r)


def _uniform_tx_mask(tx_log2):
    """VP9Filter of a superblock tiled by one transform size (mask_edges, vp9block.c:1142-1262):
    luma only; mask[0][dir][y][k] k = 0:16-wide, 1:8-wide, 2:4-wide, 3:inner 4x4 edges."""
    filt = np.zeros(192, dtype=np.uint8)
    filt[:64] = 32
    m = filt[64:].reshape(2, 2, 8, 4)
    step8 = max(1, (4 << tx_log2) // 8)
    bits = sum(1 << b for b in range(0, 8, step8))
    k = 0 if tx_log2 >= 2 else 1
    for y in range(8):
        m[0, 0, y, k] = bits                 # column edges on every 8-row
        if y % step8 == 0:
            m[0, 1, y, k] = 0xFF             # row edges
        if tx_log2 == 0:
            m[0, 0, y, 3] = 0xFF
            m[0, 1, y, 3] = 0xFF
    return filt


def test_schedule_uniform_transforms_are_shallow():
    """Uniform 8x8 / 16x16 / 32x32 areas: every column (row) op of a pass is independent, so the
    whole superblock takes 3 steps (left edge, the other column edges, the row edges); 4x4 areas
    are a true serial chain."""
    lib = model_lib()
    for tx, expect in ((1, 3), (2, 3), (3, 3)):
        r, steps = _check(lib, _uniform_tx_mask(tx), 0, 0, 0, 7, 8)
        assert r == 0 and steps == expect, (tx, r, steps)
    r, steps = _check(lib, _uniform_tx_mask(0), 0, 0, 0, 7, 8)
    assert r == 0 and steps == 32
