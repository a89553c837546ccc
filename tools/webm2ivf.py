#!/usr/bin/env python3
"""Remux the VP9 track of a WebM / Matroska file into IVF (the container the parity tools read,
reference libavformat/ivfdec.c:29-100).  A minimal EBML walker: Segment -> Cluster ->
SimpleBlock / BlockGroup.Block, no lacing (video tracks written by libavformat never lace).

    tools/webm2ivf.py in.webm out.ivf
"""
from __future__ import annotations

import struct
import sys
from pathlib import Path
from typing import Iterator, List, Tuple

ID_SEGMENT, ID_CLUSTER, ID_SIMPLEBLOCK, ID_BLOCKGROUP, ID_BLOCK = 0x18538067, 0x1F43B675, 0xA3, 0xA0, 0xA1
ID_TRACKS, ID_TRACKENTRY, ID_VIDEO, ID_PIXW, ID_PIXH, ID_CODECID = 0x1654AE6B, 0xAE, 0xE0, 0xB0, 0xBA, 0x86
MASTERS = {ID_SEGMENT, ID_CLUSTER, ID_BLOCKGROUP, ID_TRACKS, ID_TRACKENTRY, ID_VIDEO}


def _vint(b: bytes, pos: int, keep_marker: bool) -> Tuple[int, int]:
    first = b[pos]
    n = 1
    while n <= 8 and not first & (0x80 >> (n - 1)):
        n += 1
    if n > 8:
        raise ValueError(f"bad EBML vint at {pos}")
    v = first if keep_marker else first & (0xFF >> n)
    for i in range(1, n):
        v = v << 8 | b[pos + i]
    if not keep_marker and v == (1 << (7 * n)) - 1:
        v = -1                                  # unknown size
    return v, pos + n


def _walk(b: bytes, pos: int, end: int) -> Iterator[Tuple[int, int, int]]:
    while pos < end:
        eid, pos = _vint(b, pos, True)
        size, pos = _vint(b, pos, False)
        stop = end if size < 0 else pos + size
        if eid in MASTERS:
            yield eid, pos, pos               # announce the master, then descend
            yield from _walk(b, pos, stop)
        else:
            yield eid, pos, stop
        pos = stop


def read_webm(path: Path) -> Tuple[int, int, List[bytes]]:
    b = Path(path).read_bytes()
    w = h = 0
    codec = b""
    frames: List[bytes] = []
    for eid, a, z in _walk(b, 0, len(b)):
        if eid == ID_PIXW:
            w = int.from_bytes(b[a:z], "big")
        elif eid == ID_PIXH:
            h = int.from_bytes(b[a:z], "big")
        elif eid == ID_CODECID:
            codec = b[a:z]
        elif eid in (ID_SIMPLEBLOCK, ID_BLOCK) and z > a:
            _, p = _vint(b, a, False)           # track number
            flags = b[p + 2]
            if flags & 0x06:
                raise ValueError("laced blocks are not supported")
            frames.append(b[p + 3:z])
    if codec != b"V_VP9":
        raise ValueError(f"not a VP9 track: {codec!r}")
    return w, h, frames


def write_ivf(path: Path, w: int, h: int, frames: List[bytes]) -> None:
    out = bytearray(b"DKIF" + struct.pack("<HH4sHHIIII", 0, 32, b"VP90", w, h, 30, 1, len(frames), 0))
    for i, f in enumerate(frames):
        out += struct.pack("<IQ", len(f), i) + f
    Path(path).write_bytes(bytes(out))


if __name__ == "__main__":
    w_, h_, fr = read_webm(Path(sys.argv[1]))
    write_ivf(Path(sys.argv[2]), w_, h_, fr)
    print(f"{w_}x{h_}, {len(fr)} packets, {sum(map(len, fr))} bytes")
