"""Shared helpers of the slot-level tests: ctypes views of the REAL reference slots
(oracle/_ref/libvp9dsp_ref.so = oracle/slot_shim.c + unmodified reference objects) and of the
restatement (oracle/_ref/libreplay_model.so), plus checkasm-style input builders
(reference tests/checkasm/vp9dsp.c:50-625)."""
import ctypes as C
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
_ref = _model = None


def ref_lib():
    global _ref
    if _ref is None:
        _ref = C.CDLL(str(ROOT / "oracle" / "_ref" / "libvp9dsp_ref.so"))
    return _ref


def model_lib():
    global _model
    if _model is None:
        _model = C.CDLL(str(ROOT / "oracle" / "_ref" / "libreplay_model.so"))
    return _model


def ptr(a: np.ndarray, byte_off: int = 0):
    return C.c_void_p(a.ctypes.data + byte_off)


def pix_dtype(bpp):
    return np.uint16 if bpp > 8 else np.uint8


def coef_dtype(bpp):
    return np.int32 if bpp > 8 else np.int16


def rand_pixels(rng, shape, bpp):
    return rng.integers(0, 1 << bpp, size=shape).astype(pix_dtype(bpp))


def itx_cases(rng, bpp, n_per=3):
    """(tx, txtp, eob, coefs[N*N]) tuples: natural, sparse and adversarial magnitudes."""
    cd = coef_dtype(bpp)
    info = np.iinfo(cd)
    out = []
    for tx in range(5):
        t = 0 if tx == 4 else tx
        N = 4 << t
        for txtp in range(4):
            for k in range(n_per):
                c = np.zeros(N * N, cd)
                kind = k % 3
                if kind == 0:      # low-frequency natural residual
                    m = min(N, 4)
                    blk = rng.integers(-(60 << (bpp - 8)), 60 << (bpp - 8), size=(m, m))
                    c.reshape(N, N)[:m, :m] = blk
                    eob = m * m
                elif kind == 1:    # dense, moderate
                    c[:] = rng.integers(-(1 << (bpp + 2)), 1 << (bpp + 2), size=N * N)
                    eob = N * N
                else:              # adversarial: full range of the storage type
                    c[:] = rng.integers(info.min, info.max, size=N * N, endpoint=True)
                    eob = N * N
                out.append((tx, txtp, eob, c))
            c = np.zeros(N * N, cd)
            c[0] = rng.integers(-(1 << (bpp + 4)), 1 << (bpp + 4))
            out.append((tx, txtp, 1, c))                       # DC only
            c = np.zeros(N * N, cd)
            c[0] = info.max if txtp & 1 else info.min
            out.append((tx, txtp, 1, c))                       # DC only, extreme
    return out


def lf_tile(rng, bpp, kind):
    """64x64 tile with the edge at (32,32); kind picks the filter class like checkasm's
    flat16 / flat8 / regular / off inputs (tests/checkasm/vp9dsp.c:393-447)."""
    mx = (1 << bpp) - 1
    if kind == 0:   # random: mostly "off"
        return rand_pixels(rng, (64, 64), bpp)
    base = rng.integers(20 << (bpp - 8), 230 << (bpp - 8))
    amp = {1: 1 << (bpp - 8), 2: 3 << (bpp - 8), 3: 12 << (bpp - 8)}[kind]
    t = base + rng.integers(-amp, amp + 1, size=(64, 64))
    if kind == 3:   # a step across both edges
        t[:, 32:] += rng.integers(-(20 << (bpp - 8)), 20 << (bpp - 8))
        t[32:, :] += rng.integers(-(20 << (bpp - 8)), 20 << (bpp - 8))
    return np.clip(t, 0, mx).astype(pix_dtype(bpp))
