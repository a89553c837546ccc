#!/usr/bin/env python3
"""Per-frame stage times of one workload (records resident in HBM): tools/prof_frames.py c3 [frames]"""
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from ffmpeg_hybrid_b200 import binding, parity, v9r  # noqa: E402
import numpy as np  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
    nfr = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    lib = binding.load()
    tmp = Path(tempfile.mkdtemp())
    ivf = parity.write_stream(bench.make_spec(wl, 0, nfr), tmp / "s.ivf")
    parity.dump_records(ivf, tmp / "s.v9r")
    raw = v9r.read_v9r(tmp / "s.v9r")
    frames = [binding.HostFrame(lib, fr) for fr in raw]
    ctx = binding.Context(0)
    handles = [ctx.stage(f) for f in frames]
    for rep in range(2):
        for h in handles:
            ctx.wait(ctx.run_staged(h))
    ctx.option("timing", 1)
    ctx.option("phase_clk", 1)
    names = None
    for i, h in enumerate(handles):
        ctx.option("reset_stats")
        ctx.wait(ctx.run_staged(h))
        st = ctx.kernel_stats()
        itx = np.frombuffer(b"".join(l.itx for l in raw[i].lanes), dtype=v9r.ITX_DTYPE)
        nlv = int(itx["level"].max()) + 1 if len(itx) else 0
        if names is None:
            names = [k for k in st]
            print("frame  w    h  n_itx  levels " + " ".join(f"{k[:10]:>10s}" for k in names))
        print(f"{i:4d} {frames[i].w:5d} {frames[i].h:4d} {len(itx):6d} {nlv:6d} " + " ".join(f"{st[k]['ms']:10.3f}" for k in names))
    pc = ctx.phase_clocks(24)
    units = sum(((f.w + 63) // 64) * ((f.h + 63) // 64) for f in frames)
    for name, v, n in (("luma", pc[:5], units), ("chroma", pc[5:10], 2 * units)):
        print(f"loop filter {name}: cycles per superblock turn (wait body, Cr, wait top, R, Ce+signal):",
              [round(x / n) for x in v], "sum", round(sum(v) / n))
    turns = max(pc[14], 1)
    print(f"intra: non-empty turns {pc[14]}, steps/turn {pc[13] / turns:.1f}, blocks/turn {pc[15] / turns:.1f}, "
          f"cycles/turn wait body {pc[10] / turns:.0f}, wait top {pc[11] / turns:.0f}, compute {pc[12] / turns:.0f} "
          f"(= {pc[12] / max(pc[13], 1):.0f} per step); [-DVP9CUDA_STEPSTATS builds only] chunks/step {pc[17] / max(pc[13], 1):.1f}, "
          f"rounds/step {pc[16] / max(pc[13], 1):.2f}, turns that wait for the row above {pc[18] / turns:.2f}")
    ctx.close()


if __name__ == "__main__":
    main()
