"""Pins the oracle restatement's arithmetic (ffmpeg_hybrid_b200/csrc/vp9_{itx,pred,lf}.h compiled
for the host in oracle/replay_model.cc) against the UNMODIFIED reference VP9DSPContext slots,
slot by slot, checkasm style (reference tests/checkasm/vp9dsp.c).  Bit-exact or fail."""
import ctypes as C

import numpy as np
import pytest

from common import (coef_dtype, itx_cases, lf_tile, model_lib, pix_dtype, ptr, rand_pixels, ref_lib)

BPPS = [8, 10, 12]


@pytest.mark.parametrize("bpp", BPPS)
def test_itxfm_add(bpp):
    rng = np.random.default_rng(100 + bpp)
    ref, mod = ref_lib(), model_lib()
    bps = 2 if bpp > 8 else 1
    for tx, txtp, eob, coefs in itx_cases(rng, bpp, n_per=6):
        dst0 = rand_pixels(rng, (32, 64), bpp)
        a, b = dst0.copy(), dst0.copy()
        ca, cb = coefs.copy(), coefs.copy()
        ref.ref_itxfm_add(bpp, tx, txtp, ptr(a), C.c_ssize_t(64 * bps), ptr(ca), eob)
        mod.model_itxfm_add(bpp, tx, txtp, ptr(b), C.c_ssize_t(64 * bps), ptr(cb), eob)
        assert np.array_equal(a, b), f"itxfm_add mismatch bpp={bpp} tx={tx} txtp={txtp} eob={eob}"
        assert not ca.any() or eob == 1, "reference slot must zero its block"


@pytest.mark.parametrize("bpp", BPPS)
def test_intra_pred(bpp):
    rng = np.random.default_rng(200 + bpp)
    ref, mod = ref_lib(), model_lib()
    bps = 2 if bpp > 8 else 1
    for tx in range(4):
        N = 4 << tx
        for mode in range(15):
            for rep in range(4):
                top = rand_pixels(rng, 96, bpp)
                left = rand_pixels(rng, 64, bpp)
                if rep == 3:                       # saturated edges (TM clipping)
                    top[:] = (1 << bpp) - 1
                    left[:] = 0 if mode != 9 else (1 << bpp) - 1
                a = np.zeros((32, 64), pix_dtype(bpp))
                b = np.zeros((32, 64), pix_dtype(bpp))
                ref.ref_intra_pred(bpp, tx, mode, ptr(a), C.c_ssize_t(64 * bps), ptr(left), ptr(top, 16 * bps))
                mod.model_intra_pred(bpp, tx, mode, ptr(b), C.c_ssize_t(64 * bps), ptr(left), ptr(top, 16 * bps))
                assert np.array_equal(a[:N, :N], b[:N, :N]), f"intra_pred mismatch bpp={bpp} tx={tx} mode={mode}"


@pytest.mark.parametrize("bpp", BPPS)
def test_mc_and_scaled_mc(bpp):
    rng = np.random.default_rng(300 + bpp)
    ref, mod = ref_lib(), model_lib()
    bps = 2 if bpp > 8 else 1
    S = 320
    for hsize, w in enumerate([64, 32, 16, 8, 4]):
        for filt in range(4):
            for avg in range(2):
                for (mx, my) in [(0, 0), (5, 0), (0, 9), (3, 14), (15, 15), (8, 1)]:
                    src = rand_pixels(rng, (S, S), bpp)
                    if (mx + my) % 3 == 0:
                        src[::2] = (1 << bpp) - 1       # provoke over/undershoot clipping
                        src[1::2] = 0
                    h = int(rng.choice([w, max(4, w // 2)]))
                    d0 = rand_pixels(rng, (64, 64), bpp)
                    a, b = d0.copy(), d0.copy()
                    off = (72 * S + 72) * bps
                    ref.ref_mc(bpp, hsize, filt, avg, ptr(a), C.c_ssize_t(64 * bps), ptr(src, off), C.c_ssize_t(S * bps), h, mx, my)
                    mod.model_mc(bpp, w, filt, avg, ptr(b), C.c_ssize_t(64 * bps), ptr(src, off), C.c_ssize_t(S * bps), h, mx, my, 0, 0)
                    assert np.array_equal(a, b), f"mc mismatch bpp={bpp} w={w} filter={filt} avg={avg} mx={mx} my={my}"
                for (mx, my, dx, dy) in [(0, 0, 32, 32), (7, 3, 24, 17), (15, 1, 8, 12), (4, 4, 16, 16), (9, 0, 1, 2)]:
                    src = rand_pixels(rng, (S, S), bpp)
                    h = int(rng.choice([w, max(4, w // 2)]))
                    d0 = rand_pixels(rng, (64, 64), bpp)
                    a, b = d0.copy(), d0.copy()
                    off = (72 * S + 72) * bps
                    ref.ref_smc(bpp, hsize, filt, avg, ptr(a), C.c_ssize_t(64 * bps), ptr(src, off), C.c_ssize_t(S * bps), h, mx, my, dx, dy)
                    mod.model_mc(bpp, w, filt, avg, ptr(b), C.c_ssize_t(64 * bps), ptr(src, off), C.c_ssize_t(S * bps), h, mx, my, dx, dy)
                    assert np.array_equal(a, b), f"smc mismatch bpp={bpp} w={w} filter={filt} avg={avg} {mx},{my},{dx},{dy}"


@pytest.mark.parametrize("bpp", BPPS)
def test_loop_filter(bpp):
    rng = np.random.default_rng(400 + bpp)
    ref, mod = ref_lib(), model_lib()
    bps = 2 if bpp > 8 else 1
    off = (32 * 64 + 32) * bps
    for kind in range(4):
        for rep in range(12):
            for wd_i, wd in enumerate([4, 8, 16]):
                for d in range(2):
                    L = int(rng.integers(1, 64))
                    E, I, H = int(rng.integers(1, 255)), int(rng.integers(1, 64)), L >> 4
                    t0 = lf_tile(rng, bpp, kind)
                    a, b = t0.copy(), t0.copy()
                    ref.ref_loop_filter_8(bpp, wd_i, d, ptr(a, off), C.c_ssize_t(64 * bps), E, I, H)
                    mod.model_loop_filter(bpp, wd, d, 8, ptr(b, off), C.c_ssize_t(64 * bps), E, I, H)
                    assert np.array_equal(a, b), f"loop_filter_8 mismatch bpp={bpp} wd={wd} dir={d} kind={kind}"
                    if wd == 16:
                        a, b = t0.copy(), t0.copy()
                        ref.ref_loop_filter_16(bpp, d, ptr(a, off), C.c_ssize_t(64 * bps), E, I, H)
                        mod.model_loop_filter(bpp, 16, d, 16, ptr(b, off), C.c_ssize_t(64 * bps), E, I, H)
                        assert np.array_equal(a, b), f"loop_filter_16 mismatch bpp={bpp} dir={d} kind={kind}"


def _shift_lines(t, d, n):
    """the tile moved by n lines against the line direction of an edge (dir 0: column edge, lines run down)"""
    return np.roll(t, -n, axis=0 if d == 0 else 1)


@pytest.mark.parametrize("bpp", BPPS)
def test_loop_filter_mix2(bpp):
    """loop_filter_mix2[wd1][wd2][dir] (vp9dsp.h:96-104): two adjacent 8-line filters of width 4 or 8 with their
    own limits in one call -- the restatement runs it as two 8-line calls"""
    rng = np.random.default_rng(450 + bpp)
    ref, mod = ref_lib(), model_lib()
    bps = 2 if bpp > 8 else 1
    off = (32 * 64 + 32) * bps
    for kind in range(4):
        for rep in range(6):
            for w1 in range(2):
                for w2 in range(2):
                    for d in range(2):
                        E = [int(rng.integers(1, 255)) for _ in range(2)]
                        I = [int(rng.integers(1, 64)) for _ in range(2)]
                        H = [int(rng.integers(0, 4)) for _ in range(2)]
                        t0 = lf_tile(rng, bpp, kind)
                        a = t0.copy()
                        ref.ref_loop_filter_mix2(bpp, w1, w2, d, ptr(a, off), C.c_ssize_t(64 * bps),
                                                 E[0] | E[1] << 8, I[0] | I[1] << 8, H[0] | H[1] << 8)
                        b0, b1 = t0.copy(), _shift_lines(t0, d, 8).copy()
                        mod.model_loop_filter(bpp, 4 << w1, d, 8, ptr(b0, off), C.c_ssize_t(64 * bps), E[0], I[0], H[0])
                        mod.model_loop_filter(bpp, 4 << w2, d, 8, ptr(b1, off), C.c_ssize_t(64 * bps), E[1], I[1], H[1])
                        sl0 = (slice(32, 40), slice(None)) if d == 0 else (slice(None), slice(32, 40))
                        sl1 = (slice(40, 48), slice(None)) if d == 0 else (slice(None), slice(40, 48))
                        assert np.array_equal(a[sl0], b0[sl0]) and np.array_equal(a[sl1], b1[sl0]), (bpp, w1, w2, d, kind)
