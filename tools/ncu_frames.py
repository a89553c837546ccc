#!/usr/bin/env python3
"""Decode the first frames of a workload once from HBM-resident records (for ncu captures):
tools/ncu_frames.py c3 [frames]"""
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from ffmpeg_hybrid_b200 import binding, parity, v9r  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
    nfr = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    lib = binding.load()
    tmp = Path(tempfile.mkdtemp())
    ivf = parity.write_stream(bench.make_spec(wl, 0, nfr), tmp / "s.ivf")
    parity.dump_records(ivf, tmp / "s.v9r")
    raw = v9r.read_v9r(tmp / "s.v9r")
    frames = [binding.HostFrame(lib, fr) for fr in raw]
    ctx = binding.Context(0)
    handles = [ctx.stage(f) for f in frames]
    for h in handles:
        ctx.wait(ctx.run_staged(h))
    print("checksum", hex(ctx.checksum(-1)))
    ctx.close()


if __name__ == "__main__":
    main()
