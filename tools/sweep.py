#!/usr/bin/env python3
"""Frames/s (records resident) of a workload for several values of one set_option key:
tools/sweep.py <workload> <frames> <key> v1 v2 ...   (e.g. c3 0 intra_poll_ns 0 50 100 200)"""
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from ffmpeg_hybrid_b200 import binding, parity  # noqa: E402


def main():
    wl, nfr, key = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    vals = [int(v) for v in sys.argv[4:]]
    lib = binding.load()
    tmp = Path(tempfile.mkdtemp())
    ivf = parity.write_stream(bench.make_spec(wl, 0, nfr), tmp / "s.ivf")
    parity.dump_records(ivf, tmp / "s.v9r")
    frames = binding.load_stream(lib, tmp / "s.v9r", pinned=True)
    shown = sum(1 for f in frames if not f.invisible)
    for v in vals:
        ctx = binding.Context(0, frames_in_flight=3)
        ctx.option(key, v)
        hs = [ctx.stage(f) for f in frames]
        for _ in range(3):
            for h in hs:
                t = ctx.run_staged(h, None)
        ctx.wait(t)
        ctx.option("mark", 0)
        steps = 5
        for _ in range(steps):
            for h in hs:
                t = ctx.run_staged(h, None)
        ctx.option("mark", 1)
        ctx.wait(t)
        ms = ctx.elapsed_ms(0, 1)
        print(f"{wl} {key}={v}: {shown * steps / (ms * 1e-3):.1f} frames/s  checksum {ctx.checksum(-1):016x}", flush=True)
        ctx.close()


if __name__ == "__main__":
    main()
