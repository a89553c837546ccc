#!/usr/bin/env python3
"""Aggregate frames/s of S independent streams decoded side by side on ONE GPU (one context and one
host thread per stream, records resident in HBM): how much of the latency-bound wavefront's idle
issue capacity concurrent streams can fill.  Information only -- bench.py's figures are one stream
per GPU.

    python tools/multi_stream.py [--workload c3] [--streams 1,2,4] [--steps 5]
"""
import argparse
import json
import sys
import tempfile
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--streams", default="1,2,4")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--frames", type=int, default=0)
    args = ap.parse_args()
    import torch
    assert torch.cuda.is_available()
    import bench
    from ffmpeg_hybrid_b200 import binding, parity
    lib = binding.load()
    tmp = Path(tempfile.mkdtemp(prefix="vp9ms_"))
    spec = bench.make_spec(args.workload, 0, args.frames)
    ivf = parity.write_stream(spec, tmp / "s.ivf")
    parity.dump_records(ivf, tmp / "s.v9r")
    frames = binding.load_stream(lib, tmp / "s.v9r", pinned=True)
    shown = sum(1 for f in frames if not f.invisible)
    res = {}
    for S in [int(x) for x in args.streams.split(",")]:
        ctxs = [binding.Context(device=0, frames_in_flight=3) for _ in range(S)]
        handles = [[c.stage(f) for f in frames] for c in ctxs]
        go = threading.Barrier(S + 1)
        done = threading.Barrier(S + 1)
        sums = [None] * S

        def worker(i):
            c, hs = ctxs[i], handles[i]
            for _ in range(3):
                t = 0
                for h in hs:
                    t = c.run_staged(h, None)
                c.wait(t)
            go.wait()
            t = 0
            for _ in range(args.steps):
                for h in hs:
                    t = c.run_staged(h, None)
            c.wait(t)
            done.wait()
            sums[i] = c.checksum(-1)

        ts = [threading.Thread(target=worker, args=(i,)) for i in range(S)]
        for t in ts:
            t.start()
        go.wait()
        t0 = time.perf_counter()
        done.wait()
        dt = time.perf_counter() - t0
        for t in ts:
            t.join()
        assert len(set(sums)) == 1, sums          # every stream ends on the same picture
        res[S] = round(S * shown * args.steps / dt, 1)
        for c in ctxs:
            c.close()
    print(json.dumps({"workload": args.workload, "frames_per_sec_by_streams_on_one_gpu": res,
                      "timing": "host clock around all streams, records resident"}))


if __name__ == "__main__":
    main()
