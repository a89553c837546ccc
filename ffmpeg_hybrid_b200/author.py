"""Syntax-driven test streams: the reference's own parser run as a WRITER (tools/vp9author.c).

A seed stream from streamgen.py supplies real frame headers; inside the tiles every symbol the parser asks for is
chosen by the authoring policy and bool-encoded under the probability the parser itself computed, so partitions,
modes, motion vectors (final vectors, against the find_ref_mvs predictor) and coefficient tokens are written in
exactly the order vp9block.c:80-923 reads them.  oracle/_ref/author/ffvp9_author is test tooling built from the
reference sources (tools/vp9author.mk); it travels to the GPU box prebuilt like the rest of oracle/_ref."""
from __future__ import annotations

import copy
import os
import subprocess
from pathlib import Path
from typing import Dict, Tuple

from . import configs, parity
from .streamgen import StreamSpec

AUTHOR_BIN = parity.ROOT / "oracle" / "_ref" / "author" / "ffvp9_author"


def author_stream(spec: StreamSpec, policy: str, out: Path) -> Path:
    """seed = `spec` with a token payload; policy e.g. "seed=5,mv_range=512,newmv=60" (tools/vp9author.c)"""
    seed = copy.deepcopy(spec)
    for f in seed.frames:
        f.bytes_per_sb = 8               # the payload is never read: the writer supplies every tile symbol
    out.parent.mkdir(parents=True, exist_ok=True)
    seed_ivf = parity.write_stream(seed, out.with_suffix(".seed.ivf"))
    p = subprocess.run([str(AUTHOR_BIN), str(seed_ivf), str(out)], capture_output=True, text=True,
                       env=dict(os.environ, VP9AUTHOR=policy), timeout=600)
    seed_ivf.unlink()
    if p.returncode:
        raise RuntimeError(f"ffvp9_author failed: {p.stderr[-400:]}")
    return out


def authored_variants() -> Dict[str, Tuple[StreamSpec, str]]:
    """(seed spec, policy) of the authored parity streams"""
    out: Dict[str, Tuple[StreamSpec, str]] = {}
    # SURVEY 8(d) C2 mix: compound prediction, all filters, motion vectors uniform in +-64 pixels incl. off-frame
    out["authored_mv64"] = (configs.c2(frames=8, w=456, h=264, seed=41), "seed=41,mv_range=512,newmv=70")
    # small vectors, 1/8-pel precision everywhere (hp stays on while |predictor| < 8 pixels): every luma phase pair
    out["authored_hp"] = (configs.c1(frames=8, seed=42), "seed=42,mv_range=60,newmv=85")
    # 10-bit with scaled references (hidden half-size frames) and long vectors: the scaled path's clamps (vp9recon.c:492-628)
    out["authored_scaled"] = (configs.c3(frames=8, w=400, h=240, seed=43), "seed=43,mv_range=1024,newmv=60")
    # vectors far beyond the frame: +-256 pixels on a 200x136 frame, 12 bit
    st = configs.small_variants()["hbd12"]
    out["authored_far"] = (st, "seed=44,mv_range=2048,newmv=80")
    # large blocks (most partition trees forced to "none") under long compound vectors: the 64- and 32-wide mc slots
    # with every phase, windows of up to 71 x 71 samples hanging over the frame's edges
    out["authored_big"] = (configs.c2(frames=6, w=712, h=456, seed=45), "seed=45,mv_range=512,newmv=75,part_none=70")
    return out
