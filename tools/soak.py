#!/usr/bin/env python3
"""Soak run: decode full-size workloads over and over from HBM-resident records, several contexts at
once, and compare the checksum of EVERY frame of every pass with the first pass (whose pictures the
GPU parity tests pin to the reference).  A race between concurrently running wavefront CTAs would
show up as a checksum that differs from one pass to the next.

    tools/soak.py [--seconds 60] [--workloads c3,c3n,c5,c2] [--contexts 2]
prints one JSON line: passes, frames checked, mismatches."""
import argparse
import json
import sys
import tempfile
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from ffmpeg_hybrid_b200 import binding, parity  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=60)
    ap.add_argument("--workloads", default="c3,c3n,c5,c2")
    ap.add_argument("--contexts", type=int, default=2, help="contexts per workload, each on its own host thread")
    args = ap.parse_args()
    lib = binding.load()
    tmp = Path(tempfile.mkdtemp(prefix="vp9soak_"))
    jobs = []
    for wl in args.workloads.split(","):
        ivf = parity.write_stream(bench.make_spec(wl, 0, 0), tmp / f"{wl}.ivf")
        parity.dump_records(ivf, tmp / f"{wl}.v9r")
        frames = binding.load_stream(lib, tmp / f"{wl}.v9r", pinned=True)
        for k in range(args.contexts):
            jobs.append((f"{wl}#{k}", frames))
    stats = {name: {"passes": 0, "frames": 0, "mismatches": 0} for name, _ in jobs}
    deadline = time.time() + args.seconds
    errors = []

    def run(name, frames):
        try:
            ctx = binding.Context(0)
            handles = [ctx.stage(f) for f in frames]
            first = None
            while time.time() < deadline or first is None:
                sums = []
                for h in handles:
                    ctx.wait(ctx.run_staged(h))
                    sums.append(ctx.checksum(-1))
                if first is None:
                    first = sums
                else:
                    bad = [i for i, (a, b) in enumerate(zip(first, sums)) if a != b]
                    stats[name]["mismatches"] += len(bad)
                    if bad and len(stats[name].setdefault("first_bad", [])) < 8:
                        stats[name]["first_bad"].append({"pass": stats[name]["passes"], "frame": bad[0], "n": len(bad)})
                stats[name]["passes"] += 1
                stats[name]["frames"] += len(sums)
            ctx.close()
        except Exception as e:  # noqa: BLE001
            errors.append(f"{name}: {e!r}")

    threads = [threading.Thread(target=run, args=j) for j in jobs]
    t0 = time.time()
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    out = {"seconds": round(time.time() - t0, 1), "contexts": len(jobs), "per_context": stats,
           "frames_checked": sum(s["frames"] for s in stats.values()),
           "mismatches": sum(s["mismatches"] for s in stats.values()), "errors": errors}
    print(json.dumps(out))
    return 1 if out["mismatches"] or errors else 0


if __name__ == "__main__":
    sys.exit(main())
