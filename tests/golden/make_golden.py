#!/usr/bin/env python3
"""Regenerates tests/golden/framemd5.json with the UNMODIFIED reference decoder.

For every stream of ffmpeg_hybrid_b200.configs (small variants + the five full-size BASELINE
configs) it stores the SHA-256 of the generated IVF bytes (pins the stream generator) and the
per-frame MD5 list printed by oracle/_ref/ffvp9_ref (= `ffmpeg -f framemd5` of the reference
software decoder, built from /root/reference by oracle/Makefile).  Run in the build container:
    python tests/golden/make_golden.py
"""
import hashlib
import json
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))

from ffmpeg_hybrid_b200 import configs, parity  # noqa: E402
from ffmpeg_hybrid_b200.streamgen import encode_stream  # noqa: E402


def main():
    specs = dict(configs.small_variants())
    specs.update({"c1": configs.c1(), "c2": configs.c2(), "c3": configs.c3(), "c4": configs.c4(),
                  "c3p": configs.c3p(), "c3n": configs.c3n()})
    specs.update({f"c5_rank{r}": configs.c5(r) for r in range(8)})
    out = {}
    tmp = Path(tempfile.mkdtemp())
    for name, spec in specs.items():
        data = encode_stream(spec)
        p = tmp / f"{name}.ivf"
        p.write_bytes(data)
        md5 = parity.ref_md5(p)
        out[name] = {"ivf_sha256": hashlib.sha256(data).hexdigest(), "ivf_bytes": len(data), "framemd5": md5}
        print(name, len(md5), "frames")
        p.unlink()
    (Path(__file__).parent / "framemd5.json").write_text(json.dumps(out, indent=1) + "\n")
    # streams with one deliberately broken frame: the pictures that still come out + the failing packet
    bad = {}
    for name, spec in configs.damaged_variants().items():
        data = encode_stream(spec)
        p = tmp / f"{name}.ivf"
        p.write_bytes(data)
        bad[name] = {"ivf_sha256": hashlib.sha256(data).hexdigest(), "ivf_bytes": len(data),
                     "output": parity.ref_md5(p, allow_errors=True)}
        p.unlink()
    (Path(__file__).parent / "damaged.json").write_text(json.dumps(bad, indent=1) + "\n")


if __name__ == "__main__":
    main()
