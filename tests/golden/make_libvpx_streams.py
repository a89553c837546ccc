#!/usr/bin/env python3
"""Independently ENCODED test streams: OpenCV's bundled libvpx (cv2.VideoWriter 'VP90', the route
SURVEY.md section 8c verified) encodes synthetic moving pictures, tools/webm2ivf.py remuxes the
WebM to IVF, and the UNMODIFIED reference decoder (oracle/_ref/ffvp9_ref) provides the per-frame
MD5 list.  Unlike the random-payload streams of streamgen.py these carry what a real encoder
chooses: large skipped inter blocks, sub-pel motion vectors found by a motion search, sparse
coefficients, rate-controlled quantisers and loop-filter levels, hidden alt-ref frames packed
into superframes, compound prediction.

The IVF files are committed (tests/golden/libvpx_*.ivf): libvpx output depends on its version and
thread count, so the tests must not re-encode.  Regenerate in the build container with
    python tests/golden/make_libvpx_streams.py
which rewrites the .ivf files and tests/golden/libvpx.json."""
import hashlib
import json
import os
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))

STREAMS = {
    # name: (width, height, frames, libvpx options through OPENCV_FFMPEG_WRITER_OPTIONS)
    "libvpx_cif": (352, 288, 24, "lag-in-frames;12|auto-alt-ref;1|cpu-used;2|threads;1|b;400000"),
    "libvpx_1080p": (1920, 1080, 8, "tile-columns;2|cpu-used;5|threads;4|b;6000000|lag-in-frames;0"),
    "libvpx_odd": (450, 270, 10, "cpu-used;3|threads;1|crf;30|b;0|lag-in-frames;8|auto-alt-ref;1"),
}


def picture(w, h, i, seed):
    """smooth moving textures + a moving inverted box + mild noise; a scene cut two thirds in"""
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    k = i if i < 16 else i + 40
    cut = 0.0 if i < 16 else 1.7
    a = 128 + 60 * np.sin((x + 3.3 * k) / 37.0 + cut) * np.cos((y - 1.7 * k) / 23.0)
    b = 128 + 50 * np.sin((x * 0.7 + y * 0.3 + 5.1 * k) / 11.0 + cut)
    c = 128 + 70 * np.cos((x - y + 2.2 * k) / 53.0 - cut)
    img = np.stack([a, b, c], -1)
    img += np.random.default_rng([seed, i]).normal(0, 3, img.shape)
    bw, bh = max(8, w // 5), max(8, h // 4)
    bx, by = (w // 7 + 5 * i) % (w - bw), (h // 9 + 3 * i) % (h - bh)
    img[by:by + bh, bx:bx + bw] = 255 - img[by:by + bh, bx:bx + bw]
    return np.clip(img, 0, 255).astype(np.uint8)


ENCODE = r"""
import os, sys, numpy as np
os.environ["OPENCV_FFMPEG_WRITER_OPTIONS"] = sys.argv[2]
sys.path.insert(0, sys.argv[4])
import cv2
from make_libvpx_streams import picture, STREAMS
w, h, n, _ = STREAMS[sys.argv[1]]
wr = cv2.VideoWriter(sys.argv[3], cv2.VideoWriter_fourcc(*"VP90"), 30, (w, h))
assert wr.isOpened()
for i in range(n):
    wr.write(picture(w, h, i, 7))
wr.release()
"""


def main():
    from ffmpeg_hybrid_b200 import parity
    import webm2ivf
    here = Path(__file__).parent
    tmp = Path(tempfile.mkdtemp())
    out = {}
    for name, (w, h, n, opts) in STREAMS.items():
        webm = tmp / f"{name}.webm"
        # a fresh interpreter per stream: OpenCV reads the option variable once
        subprocess.run([sys.executable, "-c", ENCODE, name, opts, str(webm), str(here)], check=True,
                       stderr=subprocess.DEVNULL)
        ww, hh, frames = webm2ivf.read_webm(webm)
        assert (ww, hh) == (w, h)
        ivf = here / f"{name}.ivf"
        webm2ivf.write_ivf(ivf, w, h, frames)
        md5 = parity.ref_md5(ivf)
        assert md5 == parity.ref_md5(ivf, threads=4)
        data = ivf.read_bytes()
        out[name] = {"ivf_sha256": hashlib.sha256(data).hexdigest(), "ivf_bytes": len(data), "packets": len(frames),
                     "encoder": "libvpx via cv2 %s, options %s" % (__import__("cv2").__version__, opts),
                     "framemd5": md5}
        print(name, len(frames), "packets ->", len(md5), "pictures,", len(data), "bytes")
    (here / "libvpx.json").write_text(json.dumps(out, indent=1) + "\n")


if __name__ == "__main__":
    main()
