"""Slot-level GPU parity: vp9cuda_slot_* (the batched equivalents of the VP9DSPContext slots,
same device functions as the frame kernels) against the UNMODIFIED reference slots, checkasm
style.  Bit-exact."""
import ctypes as C

import numpy as np
import pytest

from common import coef_dtype, itx_cases, lf_tile, pix_dtype, ptr, rand_pixels, ref_lib

pytestmark = pytest.mark.gpu
BPPS = [8, 10, 12]


def u8(x):
    return np.ascontiguousarray(np.asarray(x, np.uint8))


@pytest.mark.parametrize("bpp", BPPS)
def test_slot_itxfm_add(bpp):
    from ffmpeg_hybrid_b200 import binding
    lib, ref = binding.load(), ref_lib()
    rng = np.random.default_rng(500 + bpp)
    bps = 2 if bpp > 8 else 1
    cases = itx_cases(rng, bpp, n_per=6)
    n = len(cases)
    pix = rand_pixels(rng, (n, 64, 64), bpp)
    coef = np.zeros((n, 1024), coef_dtype(bpp))
    want = pix.copy()
    for i, (tx, txtp, eob, c) in enumerate(cases):
        coef[i, :c.size] = c
        cc = c.copy()
        ref.ref_itxfm_add(bpp, tx, txtp, ptr(want[i]), C.c_ssize_t(64 * bps), ptr(cc), eob)
    tx = u8([0 if c[0] == 4 else c[0] for c in cases])
    txtp = u8([(c[1] | 0x80) if c[0] == 4 else c[1] for c in cases])
    eob = np.ascontiguousarray(np.asarray([c[2] for c in cases], np.uint16))
    rc = lib.vp9cuda_slot_itxfm_add(0, bpp, n, ptr(tx), ptr(txtp), ptr(eob), ptr(pix), ptr(coef))
    assert rc == 0
    bad = [i for i in range(n) if not np.array_equal(pix[i], want[i])]
    assert not bad, f"itxfm_add differs for cases {[(cases[i][0], cases[i][1], cases[i][2]) for i in bad[:8]]}"


@pytest.mark.parametrize("bpp", BPPS)
def test_slot_intra_pred(bpp):
    from ffmpeg_hybrid_b200 import binding
    lib, ref = binding.load(), ref_lib()
    rng = np.random.default_rng(600 + bpp)
    bps = 2 if bpp > 8 else 1
    combos = [(tx, m) for tx in range(4) for m in range(15) for _ in range(3)]
    n = len(combos)
    top = rand_pixels(rng, (n, 81), bpp)
    left = rand_pixels(rng, (n, 64), bpp)
    pix = np.zeros((n, 64, 64), pix_dtype(bpp))
    want = pix.copy()
    for i, (tx, m) in enumerate(combos):
        ref.ref_intra_pred(bpp, tx, m, ptr(want[i]), C.c_ssize_t(64 * bps), ptr(left[i]), ptr(top[i], bps))
    txs, modes = u8([c[0] for c in combos]), u8([c[1] for c in combos])   # keep alive across the call
    rc = lib.vp9cuda_slot_intra_pred(0, bpp, n, ptr(txs), ptr(modes), ptr(pix), ptr(left), ptr(top))
    assert rc == 0
    for i, (tx, m) in enumerate(combos):
        N = 4 << tx
        assert np.array_equal(pix[i, :N, :N], want[i, :N, :N]), f"intra_pred differs tx={tx} mode={m}"


@pytest.mark.parametrize("bpp", BPPS)
def test_slot_mc(bpp):
    from ffmpeg_hybrid_b200 import binding
    lib, ref = binding.load(), ref_lib()
    rng = np.random.default_rng(700 + bpp)
    bps = 2 if bpp > 8 else 1
    cases = []
    for hsize, w in enumerate([64, 32, 16, 8, 4]):
        for filt in range(4):
            for avg in range(2):
                for (mx, my, dx, dy) in [(0, 0, 0, 0), (5, 0, 0, 0), (0, 9, 0, 0), (3, 14, 0, 0), (15, 15, 0, 0),
                                         (0, 0, 32, 32), (7, 3, 24, 17), (15, 1, 8, 12), (9, 0, 1, 2)]:
                    cases.append((hsize, w, int(rng.choice([w, max(4, w // 2)])), filt, avg, mx, my, dx, dy))
    n = len(cases)
    refp = rand_pixels(rng, (n, 160, 160), bpp)
    pix = rand_pixels(rng, (n, 64, 64), bpp)
    want = pix.copy()
    off = (8 * 160 + 8) * bps
    for i, (hs, w, h, filt, avg, mx, my, dx, dy) in enumerate(cases):
        if dx:
            ref.ref_smc(bpp, hs, filt, avg, ptr(want[i]), C.c_ssize_t(64 * bps), ptr(refp[i], off), C.c_ssize_t(160 * bps), h, mx, my, dx, dy)
        else:
            ref.ref_mc(bpp, hs, filt, avg, ptr(want[i]), C.c_ssize_t(64 * bps), ptr(refp[i], off), C.c_ssize_t(160 * bps), h, mx, my)
    cols = [u8([c[k] for c in cases]) for k in range(9)]                  # keep alive across the call
    rc = lib.vp9cuda_slot_mc(0, bpp, n, *[ptr(cols[k]) for k in range(1, 9)], ptr(pix), ptr(refp))
    assert rc == 0
    bad = [cases[i] for i in range(n) if not np.array_equal(pix[i], want[i])]
    assert not bad, f"mc differs for {bad[:6]}"


@pytest.mark.parametrize("bpp", BPPS)
def test_slot_loop_filter(bpp):
    from ffmpeg_hybrid_b200 import binding
    lib, ref = binding.load(), ref_lib()
    rng = np.random.default_rng(800 + bpp)
    bps = 2 if bpp > 8 else 1
    off = (32 * 64 + 32) * bps
    cases, tiles = [], []
    for kind in range(4):
        for rep in range(10):
            for wd_i, wd in enumerate([4, 8, 16]):
                for d in range(2):
                    for ln in ([8, 16] if wd == 16 else [8]):
                        L = int(rng.integers(1, 64))
                        cases.append((wd_i, wd, d, ln, int(rng.integers(1, 255)), int(rng.integers(1, 64)), L >> 4))
                        tiles.append(lf_tile(rng, bpp, kind))
    n = len(cases)
    pix = np.stack(tiles)
    want = pix.copy()
    for i, (wd_i, wd, d, ln, E, I, H) in enumerate(cases):
        if ln == 16:
            ref.ref_loop_filter_16(bpp, d, ptr(want[i], off), C.c_ssize_t(64 * bps), E, I, H)
        else:
            ref.ref_loop_filter_8(bpp, wd_i, d, ptr(want[i], off), C.c_ssize_t(64 * bps), E, I, H)
    a8 = [u8([c[k] for c in cases]) for k in (1, 2, 3)]                    # keep alive across the call
    a32 = [np.ascontiguousarray(np.asarray([c[k] for c in cases], np.int32)) for k in (4, 5, 6)]
    rc = lib.vp9cuda_slot_loop_filter(0, bpp, n, ptr(a8[0]), ptr(a8[1]), ptr(a8[2]), ptr(a32[0]), ptr(a32[1]),
                                      ptr(a32[2]), ptr(pix))
    assert rc == 0
    bad = [cases[i] for i in range(n) if not np.array_equal(pix[i], want[i])]
    assert not bad, f"loop filter differs for {bad[:6]}"


@pytest.mark.parametrize("bpp", BPPS)
def test_slot_loop_filter_mix2(bpp):
    """loop_filter_mix2[wd1][wd2][dir] (vp9dsp.h:96-104) against vp9cuda_slot_loop_filter: the second 8 lines of a
    mix2 call are the first 8 lines of the tile moved by 8 lines"""
    from ffmpeg_hybrid_b200 import binding
    lib, ref = binding.load(), ref_lib()
    rng = np.random.default_rng(850 + bpp)
    bps = 2 if bpp > 8 else 1
    off = (32 * 64 + 32) * bps
    cases, tiles, want = [], [], []
    for kind in range(4):
        for rep in range(4):
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
                        want.append(a)
                        cases.append((4 << w1, d, 8, E[0], I[0], H[0]))
                        tiles.append(t0.copy())
                        cases.append((4 << w2, d, 8, E[1], I[1], H[1]))
                        tiles.append(np.roll(t0, -8, axis=0 if d == 0 else 1).copy())
    n = len(cases)
    pix = np.stack(tiles)
    a8 = [u8([c[k] for c in cases]) for k in (0, 1, 2)]
    a32 = [np.ascontiguousarray(np.asarray([c[k] for c in cases], np.int32)) for k in (3, 4, 5)]
    rc = lib.vp9cuda_slot_loop_filter(0, bpp, n, ptr(a8[0]), ptr(a8[1]), ptr(a8[2]), ptr(a32[0]), ptr(a32[1]),
                                      ptr(a32[2]), ptr(pix))
    assert rc == 0
    for i, w in enumerate(want):
        d = cases[2 * i][1]
        sl0 = (slice(32, 40), slice(None)) if d == 0 else (slice(None), slice(32, 40))
        sl1 = (slice(40, 48), slice(None)) if d == 0 else (slice(None), slice(40, 48))
        assert np.array_equal(w[sl0], pix[2 * i][sl0]) and np.array_equal(w[sl1], pix[2 * i + 1][sl0]), cases[2 * i]
