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

  value : frames/s with every frame's records already resident in HBM (vp9cuda_run_staged),
          pictures stay in HBM.  Device time from CUDA events on the library's own stream.
  e2e   : frames/s through the C ABI with HOST buffers: per frame the records go host->device
          from pinned memory (vp9cuda_decode_frame) and every shown picture comes back
          device->host into pinned memory.
  full_decode : for information, packets in -> pictures out through the libavcodec overlay
          (host entropy decode included), the same front end the reference arm runs.
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

WORKLOADS = {
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
    threads = max(1, cores // n_streams)

    def one_pass(ttype):
        t0 = time.perf_counter()
        procs = [subprocess.Popen([str(parity.REF_BIN), "-threads", str(threads), "-thread_type", ttype, "-null", str(p)],
                                  stdout=subprocess.DEVNULL, stderr=subprocess.PIPE) for p in ivfs]
        for p in procs:
            p.communicate()
            if p.returncode:
                raise SystemExit("reference decoder failed")
        return time.perf_counter() - t0
    # both of libavcodec's thread models (frame xor slice, pthread.c:49-65); the faster one is reported
    per_type = {}
    for ttype in ("frame", "slice"):
        for _ in range(args.warmup):
            one_pass(ttype)
        per_type[ttype] = sum(one_pass(ttype) for _ in range(args.steps))
    best = min(per_type, key=per_type.get)
    total = per_type[best]
    fps = n_streams * shown * args.steps / total
    line = {
        "impl": "reference", "metric": "decoded_frames_per_sec", "value": fps, "unit": "frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u16" if args.workload == "c3" else "u8",
        "data": "synthetic",
        "config": {"workload": f"{args.workload}: {WORKLOADS[args.workload]}", "streams": n_streams,
                   "stream_content": "own seed per stream" if args.distinct_streams else "same content in every stream",
                   "frames_per_stream": shown, "decoder": "reference libavcodec vp9 (C-only DSP, no x86 asm)",
                   "threads_per_stream": threads, "thread_type": best,
                   "fps_by_thread_type": {k: n_streams * shown * args.steps / v for k, v in per_type.items()}},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads * n_streams, "kind": "reference",
                         "sample": f"{args.steps} full passes over {n_streams} stream(s) of {shown} shown frames, "
                                   f"-thread_type {best} (faster of frame / slice)"},
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
    outs = [binding.OutFrame(lib, big.w, big.h, big.bpp) for _ in range(2)]
    d2h_per_step = sum(outs[0].nbytes for f in frames if not f.invisible and (f.w, f.h) == (big.w, big.h))

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
    ctx.option("reset_stats")
    ctx.option("timing", 1)
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
    stats = ctx.kernel_stats()
    ctx.option("timing", 0)
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
            if not f.invisible and (f.w, f.h) == (big.w, big.h):
                o = outs[k & 1]
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
                "stage_ms_per_step": {k: v["ms"] / args.steps for k, v in stats.items()},
                "stage_share": {k: v["ms"] / max(sum(x["ms"] for x in stats.values()), 1e-9) for k, v in stats.items()}}
    gpu_launches = sum(v["launches"] for v in stats.values())

    total_shown = n_shown * args.steps * world
    value = total_shown / (ms_res * 1e-3)
    e2e_value = total_shown / (ms_e2e * 1e-3)

    line = {
        "metric": "decoded_frames_per_sec", "value": value, "unit": "frames/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_res / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u16" if big.bpp > 8 else "u8", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {WORKLOADS[args.workload]}", "streams_per_gpu": 1,
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

    def timed_decoder(binary, threads, ttype, env=None):
        p = subprocess.run([str(binary), "-threads", str(threads), "-thread_type", ttype, "-null", str(ivf)],
                           capture_output=True, text=True, env=env)
        m = re.search(r"fps=([0-9.]+)", p.stderr)
        if p.returncode or not m:
            raise RuntimeError(p.stderr[-200:])
        return float(m.group(1))

    if world == 1 and not args.no_concurrent:
        # For information: S closed GOPs (copies of the stream) decoded side by side on this one GPU, one
        # context and one host thread each, records resident.  The wavefront is latency-bound, so a second
        # stream fills issue slots the first leaves idle; `value` above stays one stream per GPU.
        try:
            line["concurrent_streams"] = concurrent_leg(binding, torch, frames, n_shown, local, min(args.steps, 3))
        except Exception as e:  # noqa: BLE001
            line["concurrent_streams"] = {"error": str(e)[:200]}

    if rank == 0 and not args.no_full_decode:
        # packets in -> pictures on the host through the libavcodec overlay, host entropy decode included;
        # same front end and the same "-null" (no hashing) as the CPU baseline below
        try:
            env = dict(os.environ, VP9CUDA_LIB=str(parity.CUDA_LIB), VP9CUDA_DEVICE=str(local))
            env.pop("VP9CUDA_DUMP", None)
            got = {f"{t} threads ({tt})": timed_decoder(parity.CUDA_BIN, t, tt, env)
                   for t, tt in ((1, "slice"), (min(cores, 8), "slice"))}
            bestk = max(got, key=got.get)
            line["full_decode"] = {"value": got[bestk], "unit": "frames/s", "host": bestk, "all": got,
                                   "what": "IVF packets -> libavcodec overlay (host entropy decode, one producer lane "
                                           "per tile column, one frame in flight, pinned output frames) -> vp9cuda -> "
                                           "pictures in host memory, decoder's own clock"}
        except Exception as e:  # noqa: BLE001
            line["full_decode"] = {"error": str(e)[:200]}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        got = {tt: timed_decoder(parity.REF_BIN, cores, tt) for tt in ("frame", "slice")}
        bestk = max(got, key=got.get)
        line["cpu_baseline"] = {"value": got[bestk], "unit": "frames/s", "cores": cores, "kind": "reference",
                                "sample": f"one pass over the same stream ({n_shown} shown frames), ffvp9_ref -threads {cores} "
                                          f"-thread_type {bestk} (faster of frame {got['frame']:.1f} / slice {got['slice']:.1f} "
                                          "frames/s; C-only DSP, no x86 asm)"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    for o in outs:
        o.free()
    for f in frames:
        f.free()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
