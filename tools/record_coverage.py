#!/usr/bin/env python3
"""Which VP9DSPContext slots (reference libavcodec/vp9dsp.h:40-122) does a record stream exercise?

    tools/record_coverage.py a.v9r [b.v9r ...]          # table + the slots that were never hit

Counts, from the .v9r dumps the libavcodec glue writes ($VP9CUDA_DUMP), what checkasm enumerates
(reference tests/checkasm/vp9dsp.c:50-625) as it occurs in real frames:

  itxfm_add[tx][txtp]   x eob class (dc = 1, sub = less than a quarter of the block, full)
  intra_pred[tx][mode]  mode AFTER the edge conversion of check_intra_mode (vp9recon.c:49-70),
                        x (have_left, have_top, have_right)
  mc[hsize][filter][avg][mx != 0][my != 0] for unscaled references, every (mx, my) phase 0..15,
                        smc[hsize][filter][avg] x (dx, dy) for scaled ones,
                        motion vectors that reach across each frame edge
  loop_filter_8[wd][dir], loop_filter_16[dir], loop_filter_mix2[wd1][wd2][dir] per plane, from the
                        VP9Filter masks walked by the restatement of vp9lpf.c:31-181
                        (oracle/_ref/libreplay_model.so: model_lf_decode)

This is test tooling: it reads records and calls the oracle's mask walker, never the CUDA path.
"""
from __future__ import annotations

import ctypes as C
import sys
from collections import Counter
from pathlib import Path
from typing import Dict, Iterable, List

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from ffmpeg_hybrid_b200 import v9r  # noqa: E402

BS_W = [64, 64, 32, 32, 32, 16, 16, 16, 8, 8, 8, 4, 4]      # luma prediction width per enum BlockSize
BS_H = [64, 32, 64, 32, 16, 32, 16, 8, 16, 8, 4, 8, 4]
MODES = ["VERT", "HOR", "DC", "D45", "D135", "D117", "D153", "D63", "D207", "TM", "LEFT_DC", "TOP_DC",
         "DC_128", "DC_127", "DC_129"]
# FilterMode of the records (0 smooth, 1 regular, 2 sharp, 3 bilinear) -> slot index of mc[][N_FILTERS]
# (vp9dsp.h:32-38: FILTER_8TAP_SMOOTH, _REGULAR, _SHARP, FILTER_BILINEAR)
FILTERS = ["smooth", "regular", "sharp", "bilinear"]


def mode_conv(mode: int, have_left: bool, have_top: bool) -> int:
    """mode_conv of check_intra_mode, vp9recon.c:49-70 (same rule as csrc/vp9_pred.h)."""
    if mode in (0, 3, 7):
        return mode if have_top else 13
    if mode in (1, 8):
        return mode if have_left else 14
    if mode == 2:
        return (2 if have_top else 10) if have_left else (11 if have_top else 12)
    if mode == 9:
        return (9 if have_top else 1) if have_left else (0 if have_top else 14)
    return mode


def _rounded_div(a: int, b: int) -> int:
    return (a + (b >> 1)) // b if a >= 0 else -((-a + (b >> 1)) // b)


def all_slots() -> Dict[str, List[str]]:
    """every slot of VP9DSPContext a 4:2:0 stream can reach, by family"""
    out = {"itxfm_add": [], "intra_pred": [], "mc": [], "smc": [], "loop_filter": []}
    for tx in range(3):
        for tp in range(4):
            out["itxfm_add"].append(f"itxfm_add[{tx}][{tp}]")
    out["itxfm_add"] += ["itxfm_add[3][0]", "itxfm_add[4][0]"]      # 32x32 is DCT only; [4] = lossless WHT
    for tx in range(4):
        for m in range(15):
            out["intra_pred"].append(f"intra_pred[{tx}][{MODES[m]}]")
    for hs in range(5):
        for fi in range(4):
            for avg in range(2):
                out["smc"].append(f"smc[{hs}][{FILTERS[fi]}][{avg}]")
                for mx in range(2):
                    for my in range(2):
                        out["mc"].append(f"mc[{hs}][{FILTERS[fi]}][{avg}][{mx}][{my}]")
    for d in range(2):
        out["loop_filter"].append(f"loop_filter_16[{d}]")
        for wd in range(3):
            out["loop_filter"].append(f"loop_filter_8[{wd}][{d}]")
        for a in range(2):
            for b in range(2):
                out["loop_filter"].append(f"loop_filter_mix2[{a}][{b}][{d}]")
    return out


class Coverage:
    def __init__(self) -> None:
        self.slots: Counter = Counter()
        self.itx_eob: Counter = Counter()        # (slot, class)
        self.intra_edges: Counter = Counter()    # (slot, have_left, have_top, have_right)
        self.phases: Counter = Counter()         # (plane class, mx, my)
        self.steps: Counter = Counter()          # scaled references: (dx, dy)
        self.offframe: Counter = Counter()       # side
        self.lf_plane: Counter = Counter()       # (slot, plane class)
        self._model = None

    # ---- transforms ---------------------------------------------------------------------------
    def _itx(self, tx, txtp, eob):
        for t, tp, e in zip(tx.tolist(), txtp.tolist(), eob.tolist()):
            if e == 0:
                continue
            if tp & 0x80:
                slot = "itxfm_add[4][0]"
            else:
                slot = f"itxfm_add[{t}][{0 if t == 3 else tp & 3}]"
            n = 16 << (2 * t)
            cls = "dc" if e == 1 else "sub" if e <= n // 4 else "full"
            self.slots[slot] += 1
            self.itx_eob[(slot, cls)] += 1

    # ---- one frame ----------------------------------------------------------------------------
    def add_frame(self, fr: v9r.Frame) -> None:
        h = fr.hdr
        W, H = int(h["w"]), int(h["h"])
        for lane in fr.lanes:
            tx = np.frombuffer(lane.tx, v9r.TX_DTYPE)
            self._itx(tx["tx"], tx["txtp"], tx["eob"])
            itx = np.frombuffer(lane.itx, v9r.ITX_DTYPE)
            self._itx(itx["tx"], itx["txtp"], itx["eob"])
            for r in itx:
                hl, hr, ht = bool(r["flags"] & 1), bool(r["flags"] & 2), int(r["y4"]) > 0
                slot = f"intra_pred[{int(r['tx'])}][{MODES[mode_conv(int(r['mode']), hl, ht)]}]"
                self.slots[slot] += 1
                self.intra_edges[(slot, hl, ht, hr)] += 1
            for b in np.frombuffer(lane.blk, v9r.BLK_DTYPE):
                self._inter_block(b, h, W, H)
        if int(h["lf_level"]):
            self._loop_filter(fr)

    def _inter_block(self, b, h, W, H) -> None:
        bs, fi = int(b["bs"]), int(b["filter"])
        mv = b["mv"].astype(int)
        for r in range(2 if b["comp"] else 1):
            ref = int(b["ref"][r])
            scaled = int(h["mvscale"][ref][0]) != 0
            calls = []                                         # (plane class, width, height, x, y, mvx, mvy)
            x0, y0 = int(b["col"]) * 8, int(b["row"]) * 8
            if bs > 9:
                w, hh = BS_W[bs], BS_H[bs]
                for k in range(4):
                    px, py = 4 * (k & 1), 4 * (k >> 1)
                    if (w == 8 and px) or (hh == 8 and py):     # 8x4 / 4x8: two calls, not four
                        continue
                    calls.append((0, w, hh, x0 + px, y0 + py, mv[k][r][0], mv[k][r][1]))
                ax = _rounded_div(int(sum(mv[k][r][0] for k in range(4))), 4)
                ay = _rounded_div(int(sum(mv[k][r][1] for k in range(4))), 4)
                calls.append((1, 4, 4, x0 // 2, y0 // 2, ax, ay))
            else:
                calls.append((0, BS_W[bs], BS_H[bs], x0, y0, mv[0][r][0], mv[0][r][1]))
                calls.append((1, BS_W[bs] // 2, BS_H[bs] // 2, x0 // 2, y0 // 2, mv[0][r][0], mv[0][r][1]))
            for ch, w, hh, x, y, mvx, mvy in calls:
                hs = {64: 0, 32: 1, 16: 2, 8: 3, 4: 4}[w]
                if scaled:
                    self.slots[f"smc[{hs}][{FILTERS[fi]}][{r}]"] += 1 + ch      # chroma: U and V
                    self.steps[(int(h["mvstep"][ref][0]), int(h["mvstep"][ref][1]))] += 1
                    continue
                mx, my = (mvx & 15, mvy & 15) if ch else ((mvx & 7) << 1, (mvy & 7) << 1)
                self.slots[f"mc[{hs}][{FILTERS[fi]}][{r}][{int(mx != 0)}][{int(my != 0)}]"] += 1 + ch
                self.phases[(ch, mx, my)] += 1
                if not ch:
                    px, py = x + (mvx >> 3), y + (mvy >> 3)
                    if px - 3 < 0: self.offframe["left"] += 1
                    if py - 3 < 0: self.offframe["top"] += 1
                    if px + w + 4 > W: self.offframe["right"] += 1
                    if py + hh + 4 > H: self.offframe["bottom"] += 1

    def _loop_filter(self, fr: v9r.Frame) -> None:
        if self._model is None:
            self._model = C.CDLL(str(ROOT / "oracle" / "_ref" / "libreplay_model.so"))
        sb_cols = (int(fr.hdr["w"]) + 63) >> 6
        cw, cl, rw, rl = (np.zeros(128, np.uint8) for _ in range(4))
        for sb in range(fr.n_sb):
            raw = fr.lf[sb * 192:(sb + 1) * 192]
            for chroma in (0, 1):
                self._model.model_lf_decode(raw, chroma, int(sb % sb_cols == 0), int(sb < sb_cols),
                                            cw.ctypes.data_as(C.c_void_p), cl.ctypes.data_as(C.c_void_p),
                                            rw.ctypes.data_as(C.c_void_p), rl.ctypes.data_as(C.c_void_p))
                col, row = cw.reshape(8, 16), rw.reshape(16, 8)
                # calls cover two adjacent 8-line groups (vp9lpf.c:53-105, 131-179)
                for d, pairs in ((0, ((col[g], col[g + 1]) for g in range(0, 8, 2))),
                                 (1, ((row[:, c], row[:, c + 1]) for c in range(0, 8, 2)))):
                    for a, b in pairs:
                        for w1, w2 in zip(a.tolist(), b.tolist()):
                            if not w1 and not w2:
                                continue
                            if w1 == 16 and w2 == 16:
                                slots = [f"loop_filter_16[{d}]"]
                            elif w1 in (4, 8) and w2 in (4, 8):
                                slots = [f"loop_filter_mix2[{int(w1 == 8)}][{int(w2 == 8)}][{d}]"]
                            else:
                                slots = [f"loop_filter_8[{ {4: 0, 8: 1, 16: 2}[w] }][{d}]" for w in (w1, w2) if w]
                            for s in slots:
                                self.slots[s] += 1
                                self.lf_plane[(s, "uv" if chroma else "y")] += 1

    def add_file(self, path: Path) -> None:
        for fr in v9r.read_v9r(path):
            self.add_frame(fr)

    # ---- reports ------------------------------------------------------------------------------
    def missing(self, at_least: int = 1) -> List[str]:
        return [s for fam in all_slots().values() for s in fam if self.slots[s] < at_least]

    def report(self) -> str:
        lines = []
        for fam, names in all_slots().items():
            hit = [n for n in names if self.slots[n]]
            lines.append(f"{fam:12s} {len(hit):3d} / {len(names):3d} slots hit, "
                         f"{sum(self.slots[n] for n in names)} calls; least: "
                         f"{min((self.slots[n], n) for n in names)}")
        lines.append("itxfm eob classes: " + ", ".join(
            f"{c} {sum(v for (s, k), v in self.itx_eob.items() if k == c)}" for c in ("dc", "sub", "full")))
        lines.append("intra edge combinations (left, top, right): " + ", ".join(
            f"{int(l)}{int(t)}{int(r)}: {sum(v for (s, a, b, c), v in self.intra_edges.items() if (a, b, c) == (l, t, r))}"
            for l in (False, True) for t in (False, True) for r in (False, True)))
        for ch, name in ((0, "luma"), (1, "chroma")):
            got = {(mx, my) for (c, mx, my) in self.phases if c == ch}
            want = 64 if ch == 0 else 256
            lines.append(f"{name} sub-pel phases (mx, my) seen: {len(got)} / {want}")
        lines.append(f"scaled steps (dx, dy): {dict(self.steps)}")
        lines.append(f"prediction windows reaching across a frame edge: {dict(self.offframe)}")
        return "\n".join(lines)


def coverage_of(paths: Iterable[Path]) -> Coverage:
    c = Coverage()
    for p in paths:
        c.add_file(Path(p))
    return c


if __name__ == "__main__":
    cov = coverage_of(sys.argv[1:])
    print(cov.report())
    miss = cov.missing(1)
    print(f"never hit ({len(miss)}):", ", ".join(miss) if miss else "-")
