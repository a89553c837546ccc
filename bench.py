#!/usr/bin/env python3
"""bench.py -- decoded frames/s of the VP9 reconstruction path on B200 (BASELINE.json metric).

A "step" is one pass of the hot path over one synthetic stream (all its coded frames, in
decode order).  Per GPU the workload is one independent stream: weak scaling, no inter-GPU
traffic, no collective on the data path (torch.distributed is only the start barrier and the
max-over-ranks of the device time).  Every rank generates and decodes its own copy of the SAME
synthetic content, so that the work per GPU is fixed as N grows (streams of other seeds differ
by up to 10 % in intra-block count and therefore in decode time: 752-832 frames/s over six
seeds of C3 on one B200); --distinct-streams gives every rank its own seed instead and the
slowest stream then sets the time.

  value : RECONSTRUCTION ONLY: frames/s with every frame's records already resident in HBM
          (vp9cuda_run_staged), pictures stay in HBM.  Device time from CUDA events on the
          library's own stream, per-stage instrumentation off (stage times come from a separate
          untimed pass).  Compare with cpu_baseline.recon_only, not with cpu_baseline.value.
  e2e   : frames/s through the C ABI with HOST buffers: per frame the records go host->device
          from pinned memory (vp9cuda_decode_frame) and EVERY shown picture comes back
          device->host into pinned memory.  Still reconstruction only (records, not packets).
  full_decode : packets in -> pictures out through the libavcodec overlay (host entropy decode
          included), the same front end and the same work the reference arm does; this is the
          number to hold against cpu_baseline.value.  Beside it the host's ceiling: the same
          decoder with the record sink pointed at /dev/null (entropy decode only).
  workloads.c3p : the same measurements on C3 authored for a parallel host (frame-parallel
          headers, 8 tile columns; configs.c3p), reference decoder timed on the same stream.
  --impl reference : the unmodified reference software decoder (oracle/_ref/ffvp9_ref, built
          from /root/reference by oracle/Makefile) on all host cores, same stream(s).

Nothing here reads /root/reference; the oracle binaries are only the checker / CPU baseline.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

LOOPS = 3   # passes over the stream per decoder run; the first one is the warm-up (tools/ivf_framemd5.c -loop)

WORKLOADS = {
    "c3n": "3840x2160 10-bit profile 2 like c3 (hidden half-size frames, scaled references, 4 tile columns, backward "
           "adaptation) with NATURAL statistics: large mostly skipped inter blocks, sparse residuals, a few per cent of "
           "intra blocks (forward probability updates in every frame header, configs.natural_stats)",
    "c3p": "3840x2160 10-bit profile 2 like c3, frame-parallel headers (no backward adaptation), 8 tile columns",
    "c1": "352x288 8-bit profile 0, 30 frames",
    "c2": "1920x1080 8-bit profile 0 IPPP, 4 tile columns, compound, switchable filters, 60 frames",
    "c3": "3840x2160 10-bit profile 2, 32x32 transforms, scaled-reference (resize) frames, 30 shown + hidden frames",
    "c4": "7680x4320 8-bit all-keyframe stress stream, 8 frames",
    "c5": "3840x2160 8-bit C2-style stream (one of 8 independent streams), 30 frames",
}


def stream_index(args, rank: int) -> int:
    """Which synthetic stream a rank decodes: 0 everywhere unless --distinct-streams / VP9BENCH_STREAM."""
    if "VP9BENCH_STREAM" in os.environ:
        return int(os.environ["VP9BENCH_STREAM"])
    return rank if args.distinct_streams else 0


def make_spec(workload: str, rank: int, frames: int):
    from ffmpeg_hybrid_b200 import configs
    if workload == "c1":
        return configs.c1(frames=frames or 30, seed=1 + 100 * rank)
    if workload == "c2":
        return configs.c2(frames=frames or 60, seed=2 + 100 * rank)
    if workload == "c3":
        return configs.c3(frames=frames or 30, seed=3 + 100 * rank)
    if workload == "c3p":
        return configs.c3p(frames=frames or 30, seed=3 + 100 * rank)
    if workload == "c3n":
        return configs.c3n(frames=frames or 30, seed=3 + 100 * rank)
    if workload == "c4":
        return configs.c4(frames=frames or 8, seed=4 + 100 * rank)
    if workload == "c5":
        return configs.c5(rank, frames=frames or 30)
    raise SystemExit(f"unknown workload {workload}")


class ClockSampler(threading.Thread):
    """nvidia-smi SM clock + throttle reasons while the timed region runs."""

    def __init__(self, device: int):
        super().__init__(daemon=True)
        self.device = device
        self.stop_flag = threading.Event()
        self.sm, self.reasons, self.sm_max = [], set(), None

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device), f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                self.sm.append(float(f[0]))
                self.sm_max = float(f[1])
                for n, v in zip(names, f[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:  # noqa: BLE001
                pass
            self.stop_flag.wait(0.2)

    def result(self):
        self.stop_flag.set()
        self.join(timeout=3)
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons)}


def concurrent_leg(binding, torch, frames, n_shown, device, steps):
    """frames/s of 2 and 4 copies of the stream decoded side by side on one GPU (records resident)."""
    out = {}
    for S in (2, 4):
        ctxs = [binding.Context(device=device, frames_in_flight=3) for _ in range(S)]
        handles = [[c.stage(f) for f in frames] for c in ctxs]
        go, done = threading.Barrier(S + 1), threading.Barrier(S + 1)
        sums, errs = [None] * S, []

        def worker(i):
            try:
                c, hs = ctxs[i], handles[i]
                t = 0
                for _ in range(2):
                    for h in hs:
                        t = c.run_staged(h, None)
                c.wait(t)
                go.wait()
                for _ in range(steps):
                    for h in hs:
                        t = c.run_staged(h, None)
                c.wait(t)
                sums[i] = c.checksum(-1)
            except Exception as e:  # noqa: BLE001
                errs.append(repr(e))
                go.abort()
            finally:
                try:
                    done.wait()
                except threading.BrokenBarrierError:
                    pass

        ts = [threading.Thread(target=worker, args=(i,)) for i in range(S)]
        for t in ts:
            t.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        try:
            torch.cuda.synchronize()
            go.wait()                   # the workers have finished (and waited for) their warm-up passes
            e0.record()
            done.wait()
            e1.record()
            torch.cuda.synchronize()
        except threading.BrokenBarrierError:
            done.abort()
        for t in ts:
            t.join()
        for c in ctxs:
            c.close()
        if errs or len(set(sums)) != 1:
            raise RuntimeError(f"{S} streams: {errs or sums}")
        out[f"{S}_streams"] = S * n_shown * steps / (e0.elapsed_time(e1) * 1e-3)
    out.update(unit="frames/s (aggregate, one GPU)", steps=steps,
               what="copies of the stream (closed GOPs) decoded side by side, one context + host thread each, "
                    "records resident, CUDA events around all of them")
    return out


def run_decoders(binary, ivfs, threads, ttype, loops=LOOPS, env=None):
    """Decode every stream of `ivfs` at the same time, one process each (`-null`: no hashing); returns
    aggregate frames/s from the decoders' own steady-state clocks (tools/ivf_framemd5.c -loop: the first
    pass over a stream is the warm-up)."""
    procs = [subprocess.Popen([str(binary), "-threads", str(threads), "-thread_type", ttype, "-null", "-loop", str(loops),
                               str(p)], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, env=env)
             for p in ivfs]
    frames, secs = 0, 0.0
    for p in procs:
        _, err = p.communicate()
        m = re.search(r"frames=(\d+) seconds=([0-9.]+)", err)
        if p.returncode or not m:
            raise RuntimeError(f"{Path(binary).name} failed: {err[-300:]}")
        frames += int(m.group(1))
        secs = max(secs, float(m.group(2)))
    return frames / secs


def reference_numbers(ivfs, cores, loops=LOOPS):
    """the reference decoder on `cores` host threads over the given streams (side by side when there are
    several): both of libavcodec's thread models (frame xor slice, pthread.c:49-65), the faster one counts"""
    from ffmpeg_hybrid_b200 import parity
    threads = max(1, cores // len(ivfs))
    got = {tt: run_decoders(parity.REF_BIN, ivfs, threads, tt, loops) for tt in ("frame", "slice")}
    best = max(got, key=got.get)
    return got[best], best, threads, got


def run_reference(args, rank, world):
    """Reference arm: ffvp9_ref on the host cores, rank 0 only."""
    if rank != 0:
        return 0
    from ffmpeg_hybrid_b200 import parity
    cores = os.cpu_count() or 1
    n_streams = max(1, args.gpus)
    tmp = Path(tempfile.mkdtemp(prefix="vp9ref_"))
    ivfs, shown = [], 0
    for r in range(n_streams):
        spec = make_spec(args.workload, stream_index(args, r), args.frames)
        shown = sum(1 for f in spec.frames if f.show)
        ivfs.append(parity.write_stream(spec, tmp / f"s{r}.ivf"))
    # a "step" = one pass over the stream(s); warm-up passes are the -loop warm-up of every run
    loops = max(2, args.steps + 1)
    t0 = time.perf_counter()
    fps, best, threads, per_type = reference_numbers(ivfs, cores, loops)
    wall = time.perf_counter() - t0
    extra = {}
    if n_streams == 1 and cores >= 8:
        # one stream cannot use more cores than it has tile columns / frames in flight: the same cores
        # decoding cores/4 copies of the stream side by side (4 threads each) is the CPU's best figure
        k = cores // 4
        copies = [ivfs[0]] * k
        try:
            m_fps, m_best, m_thr, m_all = reference_numbers(copies, cores, 2)
            extra = {"multi_stream": {"value": m_fps, "unit": "frames/s (aggregate)", "streams": k,
                                      "threads_per_stream": m_thr, "thread_type": m_best,
                                      "what": "the same stream decoded side by side by that many reference processes"}}
        except Exception as e:  # noqa: BLE001
            extra = {"multi_stream": {"error": str(e)[:200]}}
    line = {
        "impl": "reference", "metric": "decoded_frames_per_sec", "value": fps, "unit": "frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * n_streams * shown / fps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u16" if args.workload in ("c3", "c3p", "c3n") else "u8",
        "data": "synthetic",
        "config": {"workload": f"{args.workload}: {WORKLOADS[args.workload]}", "streams": n_streams,
                   "stream_content": "own seed per stream" if args.distinct_streams else "same content in every stream",
                   "frames_per_stream": shown, "decoder": "reference libavcodec vp9 (C-only DSP, no x86 asm): "
                   "packets -> pictures, entropy decode + reconstruction",
                   "threads_per_stream": threads, "thread_type": best, "fps_by_thread_type": per_type,
                   "passes_per_run": loops, "wall_s": wall},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads * n_streams, "kind": "reference",
                         "sample": f"{loops - 1} timed passes (after one warm-up pass) over {n_streams} stream(s) of {shown} "
                                   f"shown frames, -thread_type {best} (faster of frame / slice)", **extra},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="vp9cuda", choices=["vp9cuda", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--frames", type=int, default=0, help="shown frames per stream (0 = config default)")
    ap.add_argument("--distinct-streams", action="store_true",
                    help="every rank decodes a stream of its own seed (default: same content on every rank)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-concurrent", action="store_true", help="skip the side-by-side streams leg")
    ap.add_argument("--no-full-decode", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "vp9cuda" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the vp9cuda back end has no CPU path")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner on stdout when the first communicator comes up: send it to
        # stderr so that stdout carries the one JSON line only
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    # keep this rank's host threads and its pinned buffers on the CPU cores next to its GPU
    affinity = None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(torch.cuda.current_device() if "CUDA_VISIBLE_DEVICES" not in os.environ
                                              else int(os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local]))
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        ideal = {64 * i + b for i, w in enumerate(words) for b in range(64) if w >> b & 1}
        mine = sorted(ideal & os.sched_getaffinity(0))
        if world > 1 and mine:
            os.sched_setaffinity(0, mine)
            affinity = f"{len(mine)} cores near the GPU (nvmlDeviceGetCpuAffinity)"
    except Exception:  # noqa: BLE001
        pass

    from ffmpeg_hybrid_b200 import binding, parity
    lib = binding.load()

    # ---- workload: synthetic stream -> (host entropy decode, untimed) record stream ----------
    tmp = Path(tempfile.mkdtemp(prefix=f"vp9bench_r{rank}_"))
    # VP9BENCH_STREAM=k: decode stream k on this rank (to tell stream content from rank effects)
    spec = make_spec(args.workload, stream_index(args, rank), args.frames)
    ivf = parity.write_stream(spec, tmp / "s.ivf")
    parity.dump_records(ivf, tmp / "s.v9r")
    frames = binding.load_stream(lib, tmp / "s.v9r", pinned=True)
    (tmp / "s.v9r").unlink()
    n_coded = len(frames)
    n_shown = sum(1 for f in frames if not f.invisible)
    h2d_per_step = sum(f.h2d_bytes for f in frames)

    ctx = binding.Context(device=local, frames_in_flight=3)
    big = max(frames, key=lambda f: f.w * f.h)
    # two pinned output pictures per picture size, used alternately: EVERY shown frame is copied out
    outs = {}
    for f in frames:
        if not f.invisible and (f.w, f.h) not in outs:
            outs[(f.w, f.h)] = [binding.OutFrame(lib, f.w, f.h, f.bpp) for _ in range(2)]
    d2h_per_step = sum(outs[(f.w, f.h)][0].nbytes for f in frames if not f.invisible)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if dist is None:
            return ms
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- leg 1: records resident in HBM ---------------------------------------------------------
    handles = [ctx.stage(f) for f in frames]

    def step_resident():
        t = 0
        for h in handles:
            t = ctx.run_staged(h, None)
        return t
    for _ in range(args.warmup):
        ctx.wait(step_resident())
    # stage split: one untimed pass with per-stage CUDA events on; the timed region below runs without them
    ctx.option("reset_stats")
    ctx.option("timing", 1)
    ctx.wait(step_resident())
    stats = ctx.kernel_stats()
    ctx.option("timing", 0)
    ctx.option("reset_stats")
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    ctx.option("mark", 0)
    t = 0
    for _ in range(args.steps):
        t = step_resident()
    ctx.option("mark", 1)
    ctx.wait(t)
    barrier()
    ms_res = max_over_ranks(ctx.elapsed_ms(0, 1))
    launches_timed = sum(v["launches"] for v in ctx.kernel_stats().values())
    checksum_resident = ctx.checksum(-1)

    if os.environ.get("VP9CUDA_PHASE"):
        ctx.option("phase_clk", 1)
        ctx.wait(step_resident())
        pc = ctx.phase_clocks(10)
        ctx.option("phase_clk", 0)
        print("loop-filter compute-warp cycles (wait body, Cr, wait top, R, Ce+signal) luma:", pc[:5],
              "chroma:", pc[5:], file=sys.stderr)

    # ---- leg 2: end to end through the C ABI with host buffers ---------------------------------
    def step_e2e():
        t = 0
        k = 0
        for f in frames:
            o = None
            if not f.invisible:
                o = outs[(f.w, f.h)][k & 1]
                k += 1
            t = ctx.decode(f, o)
        return t
    for _ in range(args.warmup):
        ctx.wait(step_e2e())
    barrier()
    w0 = time.perf_counter()
    ctx.option("mark", 2)
    for _ in range(args.steps):
        t = step_e2e()
    ctx.option("mark", 3)
    ctx.wait(t)
    wall_e2e = time.perf_counter() - w0
    barrier()
    ms_e2e = max_over_ranks(max(ctx.elapsed_ms(2, 3), 1e3 * wall_e2e))
    checksum_e2e = ctx.checksum(-1)
    clocks = sampler.result()
    if checksum_e2e != checksum_resident:
        raise SystemExit("resident and end-to-end legs produced different pictures")

    # ---- roofline of the dominant kernel (per-stage CUDA events of the timed resident leg) -----
    peaks_path = ROOT / "MEASURED_PEAKS.json"
    peak, peak_src = 6554.6, "fallback (SURVEY.md measured copy bandwidth)"
    if peaks_path.exists():
        try:
            peak = float(json.loads(peaks_path.read_text())["hbm_gbs"])
            peak_src = "MEASURED_PEAKS.json hbm_gbs"
        except Exception:  # noqa: BLE001
            pass
    kern = {k: v for k, v in stats.items() if v["launches"] and v["ms"] > 0}
    dom = max(kern, key=lambda k: kern[k]["ms"])
    d = kern[dom]
    achieved = d["bytes"] / (d["ms"] * 1e-3) / 1e9
    # measured DRAM traffic of that kernel per launch, from the committed `ncu --set full` capture
    traffic, traffic_src = None, None
    tpath = ROOT / "profiles" / f"ncu_traffic_{args.workload}.json"
    if tpath.exists():
        try:
            t = json.loads(tpath.read_text()).get(dom)
            if t:
                traffic, traffic_src = t["dram_bytes_per_launch"], t["source"]
        except Exception:  # noqa: BLE001
            pass
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "note": "wavefront kernel: bound by the dependency chain of the intra / loop-filter wavefronts, not "
                        "by bandwidth (DESIGN.md section 4); stages on the prep stream (cellmap, intra_sort, lf_ops) "
                        "overlap the compute stream, so stage times sum to more than the step",
                "avg_launch_ms": d["ms"] / d["launches"],
                "algorithmic_bytes_per_launch": d["bytes"] / d["launches"],
                "stage_ms_per_step": {k: v["ms"] for k, v in stats.items()},
                "stage_timing": "CUDA events around every stage on the stream it is launched on, in a separate pass over the "
                                "same step right before the timed region (the timed region itself runs without them)",
                "stage_share": {k: v["ms"] / max(sum(x["ms"] for x in stats.values()), 1e-9) for k, v in stats.items()}}
    gpu_launches = launches_timed

    total_shown = n_shown * args.steps * world
    value = total_shown / (ms_res * 1e-3)
    e2e_value = total_shown / (ms_e2e * 1e-3)

    line = {
        "metric": "decoded_frames_per_sec", "value": value, "unit": "frames/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_res / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u16" if big.bpp > 8 else "u8", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {WORKLOADS[args.workload]}", "streams_per_gpu": 1,
                   "value_scope": "reconstruction only: records (entropy-decoded beforehand) resident in HBM -> pictures in HBM; "
                                  "hold it against cpu_baseline.recon_only. Packets -> pictures is full_decode.",
                   "stream_content": "own seed per rank" if args.distinct_streams else
                                     "every rank generates and decodes its own copy of the same content",
                   "coded_frames_per_step": n_coded, "shown_frames_per_step": n_shown,
                   "parallelism": f"{world} independent stream(s), one per GPU, no collective",
                   "l2": "inputs larger than L2: %.0f MB of records + a %d-surface reference pool per step"
                         % (h2d_per_step / 1e6, 4),
                   "timing": "CUDA events on the library stream, max over ranks",
                   "host_affinity": affinity},
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d_per_step,
                "d2h_bytes_per_step": d2h_per_step, "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": gpu_launches, "clocks": clocks, "roofline": roofline,
        "checksum": f"{checksum_resident:016x}",
    }

    cores = os.cpu_count() or 1
    share = max(1, cores // world)            # host threads this rank may use when all ranks decode at once

    if world == 1 and not args.no_concurrent:
        # For information: S closed GOPs (copies of the stream) decoded side by side on this one GPU, one
        # context and one host thread each, records resident.  The wavefront is latency-bound, so a second
        # stream fills issue slots the first leaves idle; `value` above stays one stream per GPU.
        try:
            line["concurrent_streams"] = concurrent_leg(binding, torch, frames, n_shown, local, min(args.steps, 3))
        except Exception as e:  # noqa: BLE001
            line["concurrent_streams"] = {"error": str(e)[:200]}

    def sum_over_ranks(v):
        if dist is None:
            return v
        t = torch.tensor([v], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def host_legs(stream, with_cpu):
        """packets -> pictures through the libavcodec overlay (the work the reference arm does), the host's
        ceiling (same decoder, record sink /dev/null = entropy decode only) and the reference decoder itself,
        all on the same stream, same front end, same `-null`, steady state after one warm-up pass"""
        env = dict(os.environ, VP9CUDA_LIB=str(parity.CUDA_LIB), VP9CUDA_DEVICE=str(local))
        env.pop("VP9CUDA_DUMP", None)
        null_env = dict(os.environ, VP9CUDA_DUMP="/dev/null")
        modes = [(share, "frame"), (min(share, 8), "slice")] + ([(1, "slice")] if world == 1 else [])
        out = {}
        barrier()
        got = {f"{t} threads ({tt})": sum_over_ranks(run_decoders(parity.CUDA_BIN, [stream], t, tt, LOOPS, env))
               for t, tt in modes}
        bestk = max(got, key=got.get)
        out["full_decode"] = {"value": got[bestk], "unit": "frames/s", "host": bestk, "all": got,
                              "what": "IVF packets -> libavcodec overlay (host entropy decode on frame threads = one "
                                      "frame per worker, or on tile threads = one lane per tile column; pinned output "
                                      "frames) -> vp9cuda -> pictures in host memory; decoder's own clock, %d passes after "
                                      "one warm-up pass; aggregate over %d rank(s) decoding at the same time" % (LOOPS - 1, world)}
        barrier()
        ceil = {f"{t} threads ({tt})": sum_over_ranks(run_decoders(parity.CUDA_BIN, [stream], t, tt, LOOPS, null_env))
                for t, tt in modes}
        out["entropy_only"] = {"value": ceil[bestk], "unit": "frames/s", "host": bestk, "all": ceil,
                               "what": "the same decoder with the record sink pointed at /dev/null: the host's serial "
                                       "bool-decoder work alone, i.e. the ceiling of full_decode on this host",
                               "full_decode_over_ceiling": got[bestk] / ceil[bestk]}
        if with_cpu:
            fps, best, threads, per_type = reference_numbers([stream], cores)
            one = run_decoders(parity.REF_BIN, [stream], 1, "slice", 2)
            ent1 = ceil.get("1 threads (slice)") or run_decoders(parity.CUDA_BIN, [stream], 1, "slice", 2, null_env)
            recon1 = 1.0 / max(1.0 / one - 1.0 / ent1, 1e-9)
            out["cpu_baseline"] = {
                "value": fps, "unit": "frames/s", "cores": cores, "kind": "reference",
                "sample": f"{LOOPS - 1} passes (after one warm-up pass) over the same stream ({n_shown} shown frames), ffvp9_ref "
                          f"-threads {cores} -thread_type {best} (faster of frame {per_type['frame']:.1f} / slice "
                          f"{per_type['slice']:.1f} frames/s; C-only DSP, no x86 asm); packets -> pictures",
                "one_thread": one,
                "recon_only": {"value_one_thread": recon1, "value_all_cores_ideal": recon1 * cores, "unit": "frames/s",
                               "what": "reconstruction share of the reference: 1 / (1/ref - 1/entropy_only) on one thread "
                                       "(entropy_only = the overlay decoder with the record sink at /dev/null); the all-core "
                                       "figure assumes perfect scaling over %d cores and is the like-for-like baseline for "
                                       "`value` and `e2e`" % cores,
                               "entropy_only_one_thread": ent1},
            }
            out["ratios"] = {"full_decode_over_cpu": got[bestk] / fps,
                             "value_over_cpu_recon_only_ideal": value / (recon1 * cores),
                             "e2e_over_cpu_recon_only_ideal": e2e_value / (recon1 * cores),
                             "value_over_cpu_whole_decoder": value / fps,
                             "note": "the driver's ratio divides e2e (records -> pictures) by the reference arm (packets -> "
                                     "pictures); full_decode_over_cpu is the decoder-for-decoder figure"}
        return out

    if not args.no_full_decode:
        try:
            line.update(host_legs(ivf, with_cpu=world == 1 and not args.no_cpu_baseline))
        except Exception as e:  # noqa: BLE001
            line["full_decode"] = {"error": str(e)[:300]}
        if args.workload == "c3" and world == 1:
            # the same on C3 authored for a parallel host (c3p: frame-parallel headers, 8 tile columns) and on C3 with
            # natural statistics (c3n); the headline workload stays c3
            line["workloads"] = {}
            for wl in ("c3p", "c3n"):
                try:
                    pspec = make_spec(wl, stream_index(args, rank), args.frames)
                    pivf = parity.write_stream(pspec, tmp / f"{wl}.ivf")
                    res = dict(host_legs(pivf, with_cpu=not args.no_cpu_baseline), workload=f"{wl}: " + WORKLOADS[wl])
                    for k in ("value_over_cpu_recon_only_ideal", "e2e_over_cpu_recon_only_ideal", "value_over_cpu_whole_decoder"):
                        res.get("ratios", {}).pop(k, None)
                    # reconstruction alone on this stream: records resident in HBM, same measurement as `value`
                    parity.dump_records(pivf, tmp / f"{wl}.v9r")
                    wframes = binding.load_stream(lib, tmp / f"{wl}.v9r", pinned=True)
                    (tmp / f"{wl}.v9r").unlink()
                    wctx = binding.Context(device=local, frames_in_flight=3)
                    wh = [wctx.stage(f) for f in wframes]
                    t = 0
                    for _ in range(args.warmup):
                        for h in wh:
                            t = wctx.run_staged(h, None)
                    wctx.wait(t)
                    wctx.option("mark", 0)
                    for _ in range(args.steps):
                        for h in wh:
                            t = wctx.run_staged(h, None)
                    wctx.option("mark", 1)
                    wctx.wait(t)
                    shown_w = sum(1 for f in wframes if not f.invisible)
                    res["value"] = {"value": shown_w * args.steps / (wctx.elapsed_ms(0, 1) * 1e-3), "unit": "frames/s",
                                    "what": "reconstruction only, records resident in HBM (like the headline `value`)",
                                    "record_bytes_per_step": sum(f.h2d_bytes for f in wframes)}
                    wctx.close()
                    for f in wframes:
                        f.free()
                    line["workloads"][wl] = res
                    pivf.unlink()
                except Exception as e:  # noqa: BLE001
                    line["workloads"][wl] = {"error": str(e)[:300]}
    if rank == 0:
        print(json.dumps(line), flush=True)
    for pair in outs.values():
        for o in pair:
            o.free()
    for f in frames:
        f.free()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
