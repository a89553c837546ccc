#!/usr/bin/env python3
"""Regenerates tests/golden/authored.json: the syntax-driven streams of ffmpeg_hybrid_b200.author (the reference's
parser run as a writer, tools/vp9author.c), their SHA-256 (pins the writer) and the per-frame MD5 list of the
UNMODIFIED reference decoder on them.  Run in the build container:  python tests/golden/make_authored.py"""
import hashlib
import json
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))
from ffmpeg_hybrid_b200 import author, parity  # noqa: E402


def main():
    out = {}
    tmp = Path(tempfile.mkdtemp())
    for name, (spec, policy) in author.authored_variants().items():
        ivf = author.author_stream(spec, policy, tmp / f"{name}.ivf")
        data = ivf.read_bytes()
        out[name] = {"policy": policy, "ivf_sha256": hashlib.sha256(data).hexdigest(), "ivf_bytes": len(data),
                     "framemd5": parity.ref_md5(ivf)}
        print(name, len(out[name]["framemd5"]), "frames", len(data), "bytes")
    (Path(__file__).parent / "authored.json").write_text(json.dumps(out, indent=1) + "\n")


if __name__ == "__main__":
    main()
