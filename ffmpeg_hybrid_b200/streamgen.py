"""Synthetic VP9 bitstream generator (IVF container) for the five BASELINE.json configs.

What it writes, and why it is legal for the decoder on the other side:

* the uncompressed header is written field by field in the order the reference reads it
  (libavcodec/vp9.c:527-907, grammar in cbs_vp9_syntax_template.c:268-439);
* the compressed header is produced by a real boolean *encoder* (the inverse of
  libavcodec/vpx_rac.h:84-107): marker, tx_mode, "no probability update" flags and the
  compound-prediction mode, in the read order of vp9.c:908-1110;
* every tile payload is a seeded uniform byte string whose first decoded bool (the tile
  marker, vp9.c:1345) is forced to 0.  An arithmetic decoder maps uniform bytes to symbols
  distributed exactly according to its probability model, so the payload is a *sample from
  the VP9 default/adapted model*: every partition, block size, tx size, intra mode, inter
  mode, reference, interpolation filter, motion vector and coefficient token the model gives
  non-zero probability appears, with natural-ish statistics.  The decoder never desyncs
  because there is no second party: whatever it decodes *is* the stream.  Tiles are first
  over-provisioned, then trimmed to the bytes the reference decoder actually consumed
  (see tools in tests/golden/make_golden.py).

Nothing here touches the reference tree; only numpy is needed.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

VP9_SYNCCODE = 0x498342
TX_4X4, TX_8X8, TX_16X16, TX_32X32, TX_SWITCHABLE = range(5)
FILTER_SMOOTH, FILTER_REGULAR, FILTER_SHARP, FILTER_BILINEAR, FILTER_SWITCHABLE = range(5)
# raw 2-bit header code -> FilterMode is remapped by the decoder (vp9.c:631, vp9shared.h);
# the header literal order is: 0 EIGHTTAP_SMOOTH, 1 EIGHTTAP, 2 EIGHTTAP_SHARP, 3 BILINEAR
PRED_SINGLEREF, PRED_COMPREF, PRED_SWITCHABLE = range(3)


class BitWriter:
    """MSB-first literal writer for the uncompressed header (get_bits order)."""

    def __init__(self) -> None:
        self.bits: List[int] = []

    def u(self, n: int, v: int) -> None:
        assert 0 <= v < (1 << n), (n, v)
        for i in range(n - 1, -1, -1):
            self.bits.append((v >> i) & 1)

    def s_inv(self, n: int, v: int) -> None:
        """magnitude then sign, as read by get_sbits_inv (vp9.c:379-383)."""
        self.u(n, abs(v))
        self.u(1, 1 if v < 0 else 0)

    def bytes(self) -> bytes:
        b = self.bits + [0] * (-len(self.bits) % 8)
        out = bytearray()
        for i in range(0, len(b), 8):
            v = 0
            for k in range(8):
                v = (v << 1) | b[i + k]
            out.append(v)
        return bytes(out)

    def patch_u16(self, bitpos: int, v: int) -> None:
        for i in range(16):
            self.bits[bitpos + i] = (v >> (15 - i)) & 1


class BoolEncoder:
    """Boolean arithmetic encoder matching the decoder in vpx_rac.h:84-107."""

    def __init__(self) -> None:
        self.low = 0
        self.rng = 255
        self.count = -24
        self.buf = bytearray()

    def put(self, bit: int, prob: int) -> None:
        split = 1 + (((self.rng - 1) * prob) >> 8)
        if bit:
            self.low += split
            self.rng -= split
        else:
            self.rng = split
        shift = 0
        while (self.rng << shift) < 128:
            shift += 1
        self.rng <<= shift
        self.count += shift
        if self.count >= 0:
            offset = shift - self.count
            if (self.low << (offset - 1)) & 0x80000000:
                x = len(self.buf) - 1
                while x >= 0 and self.buf[x] == 0xFF:
                    self.buf[x] = 0
                    x -= 1
                self.buf[x] += 1
            self.buf.append((self.low >> (24 - offset)) & 0xFF)
            self.low = (self.low << offset) & 0xFFFFFF
            shift = self.count
            self.count -= 8
        self.low = (self.low << shift) & 0xFFFFFFFF

    def put_eq(self, bit: int) -> None:
        self.put(bit, 128)

    def put_uint(self, n: int, v: int) -> None:
        for i in range(n - 1, -1, -1):
            self.put_eq((v >> i) & 1)

    def finish(self) -> bytes:
        for _ in range(32):
            self.put_eq(0)
        return bytes(self.buf)


@dataclass
class FrameSpec:
    """One coded frame. Field names follow VP9BitstreamHeader (vp9shared.h:103-166)."""
    keyframe: bool = False
    intraonly: bool = False
    show: bool = True
    errorres: bool = False
    width: int = 0            # 0 = inherit stream size
    height: int = 0
    refreshrefmask: int = 0x01
    refidx: Sequence[int] = (0, 0, 0)
    signbias: Sequence[int] = (0, 0, 0)
    size_from_ref: int = -1   # 0..2 = copy size from that ref, -1 = explicit size
    highprecisionmvs: bool = True
    filtermode: int = FILTER_SWITCHABLE   # header literal 0..3, or FILTER_SWITCHABLE
    refreshctx: bool = True
    parallelmode: bool = False
    framectxid: int = 0
    resetctx: int = 0
    lf_level: int = 16
    lf_sharpness: int = 0
    lf_delta_enabled: bool = True
    lf_ref_deltas: Optional[Sequence[int]] = None   # 4 values -> sent as updates
    lf_mode_deltas: Optional[Sequence[int]] = None  # 2 values
    yac_qi: int = 60
    ydc_qdelta: int = 0
    uvdc_qdelta: int = 0
    uvac_qdelta: int = 0
    log2_tile_cols: int = 0
    log2_tile_rows: int = 0
    txfmmode: int = TX_SWITCHABLE
    comppredmode: int = PRED_SWITCHABLE   # used only when sign biases differ
    bytes_per_sb: int = 1536              # over-provisioned payload per 64x64 superblock
    tile_bytes: Optional[List[int]] = None  # exact per-tile sizes (after trimming)
    show_existing: int = -1               # 0..7: header-only frame that shows reference slot N (vp9.c:1636-1653)
    # segmentation header (vp9.c:713-743); None = disabled.  Keys: "probs" (7 tree probabilities or
    # None = no map update), "pred_probs" (3, or None = no temporal prediction), "abs" and "feats"
    # (8 x (q delta | None, lf delta | None, ref 0..3 | None, skip bool), or None = no data update).
    # The segment ids themselves come out of the tile payload like every other block symbol.
    segmentation: Optional[dict] = None
    # deliberately damaged frame for the error-path tests: "marker:K" sets the marker bool of tile K
    # (vp9.c:1345 -> AVERROR_INVALIDDATA in the middle of the frame), "tilesize" writes a tile size
    # beyond the packet (vp9.c:1336-1341), "sync" breaks a key frame's sync code (vp9.c:552)
    damage: Optional[str] = None
    # forward probability updates in the compressed header (vp9.c:956-1110), coded against the context the
    # generator tracks (ProbTracker): "skip" (3 targets), "coef" ((eob, zero, one) factors f: default p ->
    # 255 - (255 - p) / f, all bands / contexts / planes), and on inter frames "intra" (4 targets) and
    # "partition" (48 targets, 64x64 level first).
    # The random payload then samples THIS model: the way to give a stream natural statistics (few intra blocks
    # in inter frames, large skipped blocks, sparse coefficients) without a syntax-level encoder.
    prob_update: Optional[dict] = None
    superframe: bool = False              # packed into one packet with the NEXT frame (superframe index,
                                          # split again by bsf/vp9_superframe_split.c:41-137)


@dataclass
class StreamSpec:
    width: int
    height: int
    profile: int = 0          # 0: 8-bit 4:2:0, 2: 10/12-bit 4:2:0, 1 / 3: the same depths with other chroma formats
    ss_h: int = 1             # chroma subsampling (profiles 1 / 3: anything but 1, 1)
    ss_v: int = 1
    bitdepth: int = 8
    seed: int = 1
    frames: List[FrameSpec] = field(default_factory=list)


def _tile_bounds(idx: int, log2_n: int, n_sb: int) -> tuple:
    """set_tile_offset (vp9.c:1244-1250), in superblocks."""
    a = min((idx * n_sb) >> log2_n, n_sb)
    b = min(((idx + 1) * n_sb) >> log2_n, n_sb)
    return a, b


def tile_col_limits(sb_cols: int) -> tuple:
    """(min_log2, max_log2) for tile columns, vp9.c:800-804."""
    mn = 0
    while sb_cols > (64 << mn):
        mn += 1
    mx = 0
    while (sb_cols >> mx) >= 4:
        mx += 1
    mx = max(0, mx - 1)
    return mn, mx


def _color_config(bw: BitWriter, st: StreamSpec) -> None:
    # read_colorspace_details (vp9.c:457-517)
    if st.profile >= 2:
        bw.u(1, 0 if st.bitdepth == 10 else 1)
    bw.u(3, 2)   # BT.709
    bw.u(1, 0)   # studio range
    if st.profile & 1:                    # profiles 1 and 3: 4:4:4, 4:2:2 or 4:4:0 (vp9.c:497-510)
        assert (st.ss_h, st.ss_v) != (1, 1), "4:2:0 is profile 0 / 2"
        bw.u(1, st.ss_h)
        bw.u(1, st.ss_v)
        bw.u(1, 0)                        # reserved
    else:
        assert (st.ss_h, st.ss_v) == (1, 1)


_INV_MAP = [7 + 13 * k for k in range(20)]
_INV_MAP += [v for v in range(1, 254) if v not in _INV_MAP] + [253]      # inv_map_table, vp9.c:401-421
_DEFAULTS: Optional[dict] = None


def default_probs() -> dict:
    """The decoder's default probability tables (VP9 specification; reference libavcodec/vp9data.c), asked
    from the decoder front end itself (`ffvp9_cuda -tables`) rather than copied."""
    global _DEFAULTS
    if _DEFAULTS is None:
        import json
        import subprocess
        from .parity import CUDA_BIN
        _DEFAULTS = json.loads(subprocess.run([str(CUDA_BIN), "-tables"], capture_output=True, text=True, check=True).stdout)
    return _DEFAULTS


def _inv_recenter_nonneg(v: int, m: int) -> int:
    if v > 2 * m:
        return v
    return m - ((v + 1) >> 1) if v & 1 else m + (v >> 1)


def _updated(d: int, p: int) -> int:
    """update_prob's result for delta index d on current probability p (vp9.c:453-454)"""
    v = _INV_MAP[d]
    return 1 + _inv_recenter_nonneg(v, p - 1) if p <= 128 else 255 - _inv_recenter_nonneg(v, 255 - p)


def _put_update(be: "BoolEncoder", p: int, target: int) -> int:
    """update flag + differential update that brings p as close to target as the code allows; returns the new value"""
    target = min(255, max(1, int(target)))
    if target == p:
        be.put(0, 252)
        return p
    d = min(range(254), key=lambda k: (abs(_updated(k, p) - target), k))
    be.put(1, 252)
    if d < 16:
        be.put_eq(0); be.put_uint(4, d)
    elif d < 32:
        be.put_eq(1); be.put_eq(0); be.put_uint(4, d - 16)
    elif d < 64:
        be.put_eq(1); be.put_eq(1); be.put_eq(0); be.put_uint(5, d - 32)
    else:
        be.put_eq(1); be.put_eq(1); be.put_eq(1)
        v = d - 64
        if v < 65:
            be.put_uint(7, v)
        else:
            be.put_uint(7, (v + 65) >> 1)
            be.put_eq((v + 65) & 1)
    return _updated(d, p)


class ProbTracker:
    """The decoder's saved probability context (context 0) for the elements the generator updates: skip, intra,
    partition and coefficient probabilities.  Differential updates (vp9.c:399-455) are relative to it, and with
    backward adaptation (vp9prob.c) it depends on what the previous frames decoded to, so it is not guessed: the
    decoder front end runs beside the generator (`ffvp9_cuda -probe`, entropy decode only, record sink /dev/null),
    is fed every finished packet and answers with the context it now holds."""

    def __init__(self) -> None:
        import os
        import subprocess
        from .parity import CUDA_BIN
        self.p = {k: list(v) for k, v in default_probs().items()}
        env = dict(os.environ, VP9CUDA_DUMP="/dev/null")
        self.proc = subprocess.Popen([str(CUDA_BIN), "-probe"], stdin=subprocess.PIPE, stdout=subprocess.PIPE, env=env)

    def feed(self, data: bytes) -> None:
        """IVF bytes: the file header (no answer), or one packet (the context after it)"""
        import json
        self.proc.stdin.write(data)
        self.proc.stdin.flush()
        if data[:4] == b"DKIF":
            return
        line = self.proc.stdout.readline()
        if not line:
            raise RuntimeError("the probing decoder gave up on the stream")
        self.p = json.loads(line)

    def close(self) -> None:
        self.proc.stdin.close()
        self.proc.wait()


def _compressed_header(fs: FrameSpec, inter: bool, allowcomp: bool, lossless: bool,
                       trk: Optional["ProbTracker"] = None) -> bytes:
    be = BoolEncoder()
    upd = fs.prob_update or {}
    assert not upd or trk is not None, "prob_update needs the tracked context (encode_stream)"
    dflt = trk.p if trk is not None else {}
    be.put(0, 128)                       # marker, vp9.c:912
    txm = TX_4X4 if lossless else fs.txfmmode
    if not lossless:                     # vp9.c:936-954
        be.put_uint(2, min(txm, 3))
        if txm >= 3:
            be.put_eq(txm - 3)
        if txm == TX_SWITCHABLE:
            for _ in range(2 + 4 + 6):
                be.put(0, 252)
    for i in range(4):                   # coef updates, vp9.c:958-992
        if "coef" in upd:
            be.put_eq(1)
            fac = upd["coef"]
            for j in range(2):
                for k in range(2):
                    for l in range(6):
                        for m in range(6):
                            if m >= 3 and l == 0:
                                break
                            base = ((((i * 2 + j) * 2 + k) * 6 + l) * 6 + m) * 3
                            for n in range(3):
                                p = dflt["coef"][base + n]
                                p0 = default_probs()["coef"][base + n]
                                _put_update(be, p, 255 - (255 - p0) / fac[n] if fac[n] != 1 else p)
        else:
            be.put_eq(0)
        if txm == i:
            break
    for i in range(3):                   # skip
        if "skip" in upd:
            _put_update(be, dflt["skip"][i], upd["skip"][i])
        else:
            be.put(0, 252)
    if inter:
        n = 21                           # mv_mode
        if fs.filtermode == FILTER_SWITCHABLE:
            n += 8
        for _ in range(n):
            be.put(0, 252)
        for i in range(4):               # intra
            if "intra" in upd:
                _put_update(be, dflt["intra"][i], upd["intra"][i])
            else:
                be.put(0, 252)
        cpm = PRED_SINGLEREF
        if allowcomp:                    # vp9.c:1016-1027
            cpm = fs.comppredmode
            be.put_eq(1 if cpm else 0)
            if cpm:
                be.put_eq(cpm - 1)
            if cpm == PRED_SWITCHABLE:
                for _ in range(5):
                    be.put(0, 252)
        n = 0
        if cpm != PRED_COMPREF:
            n += 10
        if cpm != PRED_SINGLEREF:
            n += 5
        n += 36                          # y_mode
        for _ in range(n):
            be.put(0, 252)
        for i in range(4):               # partition: coded 8x8 level first, stored 64x64 level first (vp9.c:1065-1072)
            for j in range(4):
                for k in range(3):
                    if "partition" in upd:
                        idx = ((3 - i) * 4 + j) * 3 + k
                        _put_update(be, dflt["partition"][idx], upd["partition"][idx])
                    else:
                        be.put(0, 252)
        n = 3 + 2 * 22 + 2 * 9           # motion vectors
        if fs.highprecisionmvs:
            n += 4
        for _ in range(n):
            be.put(0, 252)
    return be.finish()


def encode_frame(st: StreamSpec, fs: FrameSpec, cur_w: int, cur_h: int,
                 rng: np.random.Generator, trk: Optional["ProbTracker"] = None) -> bytes:
    """Returns one frame's bytes. cur_w/cur_h is the resolved coded size."""
    bw = BitWriter()
    bw.u(2, 2)
    bw.u(1, st.profile & 1)
    bw.u(1, st.profile >> 1)
    if st.profile == 3:
        bw.u(1, 0)                        # reserved_zero (vp9.c:535)
    if fs.show_existing >= 0:
        bw.u(1, 1)                        # show_existing_frame
        bw.u(3, fs.show_existing)
        return bw.bytes()
    bw.u(1, 0)                            # show_existing_frame
    bw.u(1, 0 if fs.keyframe else 1)
    bw.u(1, 1 if fs.show else 0)
    bw.u(1, 1 if fs.errorres else 0)
    inter = False
    allowcomp = False
    if fs.keyframe:
        bw.u(24, VP9_SYNCCODE ^ (1 if fs.damage == "sync" else 0))
        _color_config(bw, st)
        bw.u(16, cur_w - 1)
        bw.u(16, cur_h - 1)
        bw.u(1, 0)
    else:
        if not fs.show:
            bw.u(1, 1 if fs.intraonly else 0)
        else:
            assert not fs.intraonly, "intra-only frames must be hidden (vp9.c:571)"
        if not fs.errorres:
            bw.u(2, fs.resetctx)
        if fs.intraonly:
            bw.u(24, VP9_SYNCCODE)
            if st.profile >= 1:
                _color_config(bw, st)
            bw.u(8, fs.refreshrefmask)
            bw.u(16, cur_w - 1)
            bw.u(16, cur_h - 1)
            bw.u(1, 0)
        else:
            inter = True
            bw.u(8, fs.refreshrefmask)
            for i in range(3):
                bw.u(3, fs.refidx[i])
                bw.u(1, fs.signbias[i])
            if fs.size_from_ref >= 0:
                for i in range(fs.size_from_ref):
                    bw.u(1, 0)
                bw.u(1, 1)
            else:
                bw.u(3, 0)
                bw.u(16, cur_w - 1)
                bw.u(16, cur_h - 1)
            bw.u(1, 0)                    # display size
            bw.u(1, 1 if fs.highprecisionmvs else 0)
            if fs.filtermode == FILTER_SWITCHABLE:
                bw.u(1, 1)
            else:
                bw.u(1, 0)
                bw.u(2, fs.filtermode)
            sb = [0 if fs.errorres else b for b in fs.signbias]
            allowcomp = sb[0] != sb[1] or sb[0] != sb[2]
    if not fs.errorres:
        bw.u(1, 1 if fs.refreshctx else 0)
        bw.u(1, 1 if fs.parallelmode else 0)
    bw.u(2, fs.framectxid)
    # loop filter (vp9.c:669-697)
    bw.u(6, fs.lf_level)
    bw.u(3, fs.lf_sharpness)
    bw.u(1, 1 if fs.lf_delta_enabled else 0)
    if fs.lf_delta_enabled:
        upd = fs.lf_ref_deltas is not None or fs.lf_mode_deltas is not None
        bw.u(1, 1 if upd else 0)
        if upd:
            for i in range(4):
                if fs.lf_ref_deltas is not None:
                    bw.u(1, 1)
                    bw.s_inv(6, fs.lf_ref_deltas[i])
                else:
                    bw.u(1, 0)
            for i in range(2):
                if fs.lf_mode_deltas is not None:
                    bw.u(1, 1)
                    bw.s_inv(6, fs.lf_mode_deltas[i])
                else:
                    bw.u(1, 0)
    # quantisation (vp9.c:700-705)
    bw.u(8, fs.yac_qi)
    for d in (fs.ydc_qdelta, fs.uvdc_qdelta, fs.uvac_qdelta):
        if d:
            bw.u(1, 1)
            bw.s_inv(4, d)
        else:
            bw.u(1, 0)
    lossless = fs.yac_qi == 0 and not (fs.ydc_qdelta or fs.uvdc_qdelta or fs.uvac_qdelta)
    seg = fs.segmentation
    if seg is None:
        bw.u(1, 0)                        # segmentation disabled
    else:                                 # vp9.c:713-743
        bw.u(1, 1)
        probs = seg.get("probs")
        bw.u(1, 1 if probs is not None else 0)
        if probs is not None:
            for pr in probs:
                if pr is None or pr == 255:
                    bw.u(1, 0)
                else:
                    bw.u(1, 1)
                    bw.u(8, pr)
            pred = seg.get("pred_probs")
            bw.u(1, 1 if pred is not None else 0)
            if pred is not None:
                for pr in pred:
                    if pr is None or pr == 255:
                        bw.u(1, 0)
                    else:
                        bw.u(1, 1)
                        bw.u(8, pr)
        feats = seg.get("feats")
        bw.u(1, 1 if feats is not None else 0)
        if feats is not None:
            bw.u(1, 1 if seg.get("abs") else 0)
            assert len(feats) == 8
            for q, lf, ref, skip in feats:
                bw.u(1, 0 if q is None else 1)
                if q is not None:
                    bw.s_inv(8, q)
                bw.u(1, 0 if lf is None else 1)
                if lf is not None:
                    bw.s_inv(6, lf)
                bw.u(1, 0 if ref is None else 1)
                if ref is not None:
                    bw.u(2, ref)
                bw.u(1, 1 if skip else 0)
    # tiles (vp9.c:800-812)
    sb_cols = (cur_w + 63) >> 6
    sb_rows = (cur_h + 63) >> 6
    mn, mx = tile_col_limits(sb_cols)
    l2c = min(max(fs.log2_tile_cols, mn), mx)
    for _ in range(l2c - mn):
        bw.u(1, 1)
    if l2c < mx:
        bw.u(1, 0)
    l2r = fs.log2_tile_rows
    if l2r == 0:
        bw.u(1, 0)
    else:
        bw.u(1, 1)
        bw.u(1, l2r - 1)
    comp = _compressed_header(fs, inter, allowcomp, lossless, trk)
    bw.u(16, len(comp))
    out = bytearray(bw.bytes())
    out += comp
    # tile payloads (vp9.c:1325-1349)
    ntr, ntc = 1 << l2r, 1 << l2c
    k = 0
    for tr in range(ntr):
        r0, r1 = _tile_bounds(tr, l2r, sb_rows)
        for tc in range(ntc):
            c0, c1 = _tile_bounds(tc, l2c, sb_cols)
            if fs.tile_bytes is not None:
                n = fs.tile_bytes[k]
            else:
                n = 16 + (r1 - r0) * (c1 - c0) * fs.bytes_per_sb
            # always draw the over-provisioned amount so trimming keeps the same prefix
            full = 16 + (r1 - r0) * (c1 - c0) * fs.bytes_per_sb
            payload = rng.integers(0, 256, size=max(full, n), dtype=np.uint8)
            payload[0] &= 0x7F           # tile marker bool must decode to 0
            if fs.damage == f"marker:{k}":
                payload[0] |= 0x80
            payload = payload[:n]
            if tr == ntr - 1 and tc == ntc - 1:
                payload[-1] = 0          # never look like a superframe index (0b110xxxxx)
            payload = payload.tobytes()
            if not (tr == ntr - 1 and tc == ntc - 1):
                out += struct.pack(">I", len(payload) + (1 << 20 if fs.damage == "tilesize" and k == 0 else 0))
            out += payload
            k += 1
    return bytes(out)


def resolve_sizes(st: StreamSpec) -> List[tuple]:
    """Coded size of each frame, tracking the 8 reference slots like vp9.c:1687-1692."""
    slots = [(st.width, st.height)] * 8
    sizes = []
    for fs in st.frames:
        if fs.show_existing >= 0:
            sizes.append(slots[fs.show_existing])
            continue
        if fs.keyframe or fs.intraonly:
            w, h = fs.width or st.width, fs.height or st.height
        elif fs.size_from_ref >= 0:
            w, h = slots[fs.refidx[fs.size_from_ref]]
        else:
            w, h = fs.width or st.width, fs.height or st.height
        sizes.append((w, h))
        mask = 0xFF if fs.keyframe else fs.refreshrefmask
        for i in range(8):
            if mask & (1 << i):
                slots[i] = (w, h)
    return sizes


def tile_counts(st: StreamSpec) -> List[int]:
    out = []
    for fs, (w, h) in zip(st.frames, resolve_sizes(st)):
        if fs.show_existing >= 0:
            out.append(0)
            continue
        mn, mx = tile_col_limits((w + 63) >> 6)
        l2c = min(max(fs.log2_tile_cols, mn), mx)
        out.append((1 << l2c) * (1 << fs.log2_tile_rows))
    return out


def encode_stream(st: StreamSpec) -> bytes:
    sizes = resolve_sizes(st)
    out = bytearray()
    out += b"DKIF" + struct.pack("<HH4sHHIIII", 0, 32, b"VP90", st.width, st.height,
                                 30, 1, len(st.frames), 0)
    group: List[bytes] = []
    trk = ProbTracker() if any(f.prob_update for f in st.frames) else None
    if trk is not None:
        trk.feed(bytes(out))
    for i, (fs, (w, h)) in enumerate(zip(st.frames, sizes)):
        rng = np.random.default_rng([st.seed, i])
        assert not (fs.prob_update and group), "no probability updates inside a superframe (the context is probed per packet)"
        group.append(encode_frame(st, fs, w, h, rng, trk))
        if fs.superframe and i + 1 < len(st.frames) and len(group) < 8:
            continue
        if len(group) == 1:
            pkt = group[0]
        else:
            # superframe index (VP9 bitstream spec annex B): marker, frame sizes, marker
            mag = max(1, max((len(g).bit_length() + 7) // 8 for g in group))
            marker = bytes([0xC0 | (mag - 1) << 3 | (len(group) - 1)])
            pkt = b"".join(group) + marker + b"".join(len(g).to_bytes(mag, "little") for g in group) + marker
        rec = struct.pack("<IQ", len(pkt), i) + pkt
        out += rec
        if trk is not None:
            trk.feed(rec)
        group = []
    if trk is not None:
        trk.close()
    return bytes(out)
