"""Reader for the .v9r record-stream dumps written by the libavcodec glue ($VP9CUDA_DUMP).

Layout (ffmpeg_hybrid_b200/host/vp9cuda_glue.c, dump sink): "V9R3", 8 x u32 struct sizes, then per
frame "FRAM", Vp9CudaFrameHeader, n_sb, per lane {Vp9CudaLaneCounts, coef_base, blk[], tx[], itx[],
coef[] (Vp9CudaCoef16 / Vp9CudaCoef32 entries)}, sb_index[n_sb], lf[n_sb].  Struct layouts mirror
include/vp9cuda.h.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from pathlib import Path
from typing import List

import numpy as np

MAX_LANES = 64
HDR_DTYPE = np.dtype([
    ("w", "<u4"), ("h", "<u4"), ("bpp", "u1"), ("ss_h", "u1"), ("ss_v", "u1"),
    ("keyframe", "u1"), ("intraonly", "u1"), ("lossless", "u1"), ("invisible", "u1"),
    ("lf_level", "u1"), ("refreshrefmask", "u1"), ("ref_slot", "u1", (3,)),
    ("mvscale", "<u2", (3, 2)), ("mvstep", "u1", (3, 2)), ("n_lanes", "u1"), ("pad", "u1", (3,)),
    ("lane_sbs", "<u4", (MAX_LANES,)), ("lim_lut", "u1", (64,)), ("mblim_lut", "u1", (64,)),
], align=True)
COUNTS_DTYPE = np.dtype([("blk", "<u4"), ("tx", "<u4"), ("itx", "<u4"), ("coef", "<u4")])
BLK_SIZE, TX_SIZE, ITX_SIZE, SBI_SIZE, LF_SIZE = 48, 16, 20, 16, 192
BLK_DTYPE = np.dtype([("col", "<u2"), ("row", "<u2"), ("bs", "u1"), ("filter", "u1"), ("comp", "u1"), ("ref", "u1", (2,)),
                      ("pad", "u1", (3,)), ("mv", "<i2", (4, 2, 2)), ("reserved", "<u4")])
TX_DTYPE = np.dtype([("x4", "<u2"), ("y4", "<u2"), ("eob", "<u2"), ("ncoef", "<u2"), ("plane", "u1"), ("tx", "u1"),
                     ("txtp", "u1"), ("pad", "u1"), ("coef_off", "<u4")])
ITX_DTYPE = np.dtype([("x4", "<u2"), ("y4", "<u2"), ("eob", "<u2"), ("level", "<u2"), ("plane", "u1"), ("tx", "u1"),
                      ("mode", "u1"), ("txtp", "u1"), ("flags", "u1"), ("pad", "u1"), ("ncoef", "<u2"), ("coef_off", "<u4")])
SBI_DTYPE = np.dtype([("itx_start", "<u4"), ("tx_start", "<u4"), ("itx_count", "<u2"), ("tx_count", "<u2"),
                      ("lane", "u1"), ("pad", "u1", (3,))])
COEF16_DTYPE = np.dtype([("val", "<i2"), ("pos", "<u2")])
COEF32_DTYPE = np.dtype([("val", "<i4"), ("pos", "<u4")])
assert (BLK_DTYPE.itemsize, TX_DTYPE.itemsize, ITX_DTYPE.itemsize, SBI_DTYPE.itemsize) == (48, 16, 20, 16)


@dataclass
class Lane:
    counts: np.ndarray
    coef_base: int
    blk: bytes
    tx: bytes
    itx: bytes
    coef: bytes


@dataclass
class Frame:
    hdr: np.ndarray            # 0-d structured array (HDR_DTYPE)
    n_sb: int
    lanes: List[Lane] = field(default_factory=list)
    sb_index: bytes = b""
    lf: bytes = b""

    @property
    def record_bytes(self) -> int:
        n = len(self.sb_index) + len(self.lf)
        for l in self.lanes:
            n += len(l.blk) + len(l.tx) + len(l.itx) + len(l.coef)
        return n


def read_v9r(path: Path) -> List[Frame]:
    data = Path(path).read_bytes()
    if data[:4] != b"V9R3":
        raise ValueError("not a V9R3 file")
    sizes = struct.unpack("<8I", data[4:36])
    if sizes[0] != HDR_DTYPE.itemsize or sizes[1] != COUNTS_DTYPE.itemsize or sizes[2:7] != (
            BLK_SIZE, TX_SIZE, ITX_SIZE, SBI_SIZE, LF_SIZE):
        raise ValueError(f"struct sizes differ from include/vp9cuda.h: {sizes} vs {HDR_DTYPE.itemsize}")
    pos = 36
    frames: List[Frame] = []
    while pos < len(data):
        if data[pos:pos + 4] != b"FRAM":
            raise ValueError("bad frame tag")
        pos += 4
        hdr = np.frombuffer(data, HDR_DTYPE, 1, pos)[0]
        pos += HDR_DTYPE.itemsize
        (n_sb,) = struct.unpack_from("<I", data, pos)
        pos += 4
        fr = Frame(hdr=hdr, n_sb=n_sb)
        ceb = COEF32_DTYPE.itemsize if hdr["bpp"] > 8 else COEF16_DTYPE.itemsize
        for _ in range(int(hdr["n_lanes"])):
            cnt = np.frombuffer(data, COUNTS_DTYPE, 1, pos)[0]
            pos += COUNTS_DTYPE.itemsize
            (coef_base,) = struct.unpack_from("<I", data, pos)
            pos += 4

            def take(n):
                nonlocal pos
                b = data[pos:pos + n]
                pos += n
                return b
            blk = take(int(cnt["blk"]) * BLK_SIZE)
            tx = take(int(cnt["tx"]) * TX_SIZE)
            itx = take(int(cnt["itx"]) * ITX_SIZE)
            coef = take(int(cnt["coef"]) * ceb)
            fr.lanes.append(Lane(cnt, coef_base, blk, tx, itx, coef))
        fr.sb_index = data[pos:pos + n_sb * SBI_SIZE]
        pos += n_sb * SBI_SIZE
        fr.lf = data[pos:pos + n_sb * LF_SIZE]
        pos += n_sb * LF_SIZE
        frames.append(fr)
    return frames
