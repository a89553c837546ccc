"""GPU parity tests proper: everything goes through the C ABI of libvp9cuda.so, either via the
libavcodec overlay front end (ffvp9_cuda, the drop-in) or via ctypes with record dumps.
Bit-exact against the reference decoder (oracle/_ref/ffvp9_ref) and the committed golden lists."""
import hashlib
import json
from pathlib import Path

import pytest

from ffmpeg_hybrid_b200 import configs, parity

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
GOLDEN = json.loads((ROOT / "tests" / "golden" / "framemd5.json").read_text())
SMALL = sorted(configs.small_variants())


@pytest.fixture(scope="module", autouse=True)
def need_gpu():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from ffmpeg_hybrid_b200 import binding
    binding.load()


@pytest.mark.parametrize("name", SMALL)
def test_small_variants_bit_exact(name, workdir):
    ivf = parity.write_stream(configs.small_variants()[name], workdir / f"{name}.ivf")
    got = parity.cuda_md5(ivf)
    assert got == GOLDEN[name]["framemd5"], parity.compare(GOLDEN[name]["framemd5"], got)[1]
    assert got == parity.ref_md5(ivf)


@pytest.mark.parametrize("name,full", [("c1", "c1"), ("c2", "c2"), ("c3", "c3"), ("c4", "c4"), ("c5_rank0", "c5")])
def test_baseline_configs_bit_exact(name, full, workdir):
    spec = {"c1": configs.c1, "c2": configs.c2, "c3": configs.c3, "c4": configs.c4, "c5": lambda: configs.c5(0)}[full]()
    ivf = parity.write_stream(spec, workdir / f"{name}.ivf")
    got = parity.cuda_md5(ivf)
    assert got == GOLDEN[name]["framemd5"], parity.compare(GOLDEN[name]["framemd5"], got)[1]
    ivf.unlink()


@pytest.mark.parametrize("name,threads", [("c2_small", 4), ("manytiles", 16), ("segment", 4), ("p1_444", 2)])
def test_tile_threaded_host_is_bit_exact(name, threads, workdir):
    """slice threading: one producer lane per tile column (vp9.c:1442-1520)"""
    ivf = parity.write_stream(configs.small_variants()[name], workdir / "mt.ivf")
    assert parity.cuda_md5(ivf, threads=threads, thread_type="slice") == GOLDEN[name]["framemd5"]


def _md5_of(out, w, h, bpp):
    fmt = {8: "yuv420p", 10: "yuv420p10le", 12: "yuv420p12le"}[bpp]
    return f"{w} {h} {fmt} {hashlib.md5(out.bytes()).hexdigest()}"


@pytest.mark.parametrize("name", ["c1_small", "c3_small", "hbd12", "intraonly"])
def test_ctypes_record_path_and_resident_replay(name, workdir):
    """vp9cuda_decode_frame (host records -> pictures on host) and vp9cuda_stage_frame /
    vp9cuda_run_staged (records resident in HBM) give the same pictures as the reference."""
    from ffmpeg_hybrid_b200 import binding
    lib = binding.load()
    ivf = parity.write_stream(configs.small_variants()[name], workdir / f"{name}_r.ivf")
    parity.dump_records(ivf, workdir / f"{name}.v9r")
    want = [" ".join(l.split()[1:]) for l in GOLDEN[name]["framemd5"]]
    for pinned in (True, False):
        frames = binding.load_stream(lib, workdir / f"{name}.v9r", pinned=pinned)
        ctx = binding.Context(0)
        got, sums = [], []
        for f in frames:
            out = binding.OutFrame(lib, f.w, f.h, f.bpp)
            ctx.wait(ctx.decode(f, out))
            if not f.invisible:
                got.append(_md5_of(out, f.w, f.h, f.bpp))
            sums.append(ctx.checksum(-1))
            out.free()
        assert got == want
        # resident replay, twice (reference slots must cycle identically)
        handles = [ctx.stage(f) for f in frames]
        for _ in range(2):
            sums2 = []
            for h in handles:
                ctx.wait(ctx.run_staged(h))
                sums2.append(ctx.checksum(-1))
            assert sums2 == sums
        ctx.close()
        for f in frames:
            f.free()


def test_streams_decoded_side_by_side_on_one_gpu(workdir):
    """several streams in one process, one context each, driven from their own host threads at the same
    time: the wavefront kernels of different contexts share the GPU -- a CTA only
    ever waits for lower-numbered CTAs of its own launch, so they cannot block each other -- and every
    stream still gives the reference's pictures"""
    import threading
    from ffmpeg_hybrid_b200 import binding
    lib = binding.load()
    names = ["c3_small", "c2_small", "c4_small", "hbd12", "segment", "scale_mix"]
    jobs = {}
    for name in names:
        ivf = parity.write_stream(configs.small_variants()[name], workdir / f"{name}_mp.ivf")
        parity.dump_records(ivf, workdir / f"{name}_mp.v9r")
        jobs[name] = binding.load_stream(lib, workdir / f"{name}_mp.v9r", pinned=True)
    got = {n: [] for n in names}
    errs = []
    start = threading.Barrier(len(names))

    def worker(name):
        try:
            ctx = binding.Context(0)
            frames = jobs[name]
            start.wait()
            for _ in range(3):                       # three passes: the streams drift against each other
                pics = []
                for f in frames:
                    out = binding.OutFrame(lib, f.w, f.h, f.bpp)
                    ctx.wait(ctx.decode(f, out))
                    if not f.invisible:
                        pics.append(_md5_of(out, f.w, f.h, f.bpp))
                    out.free()
                got[name].append(pics)
            ctx.close()
        except Exception as e:  # noqa: BLE001
            errs.append((name, repr(e)))

    ts = [threading.Thread(target=worker, args=(n,)) for n in names]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    for name in names:
        want = [" ".join(l.split()[1:]) for l in GOLDEN[name]["framemd5"]]
        assert got[name] == [want] * 3, name
    for frames in jobs.values():
        for f in frames:
            f.free()


DAMAGED = json.loads((ROOT / "tests" / "golden" / "damaged.json").read_text())


@pytest.mark.parametrize("threads", [1, 4])
@pytest.mark.parametrize("name", sorted(configs.damaged_variants()))
def test_damaged_streams_fail_like_the_reference(name, threads, workdir):
    """a broken frame (bad sync code, tile size beyond the packet, tile marker set in the first / a later
    tile column, i.e. after begin_frame and after other tiles were submitted) is reported for the same
    packet, the frame in flight is abandoned, the reference slots stay as they were and the following
    packets decode to the reference's pictures"""
    ivf = parity.write_stream(configs.damaged_variants()[name], workdir / f"{name}.ivf")
    out = parity.cuda_md5(ivf, threads=threads, thread_type="slice" if threads > 1 else None, allow_errors=True)
    assert out == DAMAGED[name]["output"]


def test_errors_are_reported_not_hidden():
    from ffmpeg_hybrid_b200 import binding
    import ctypes as C
    lib = binding.load()
    ctx = C.c_void_p()
    assert lib.vp9cuda_create(C.byref(ctx), 99, 0, 0, 12, 2) < 0          # no such device
    c = binding.Context(0)
    assert lib.vp9cuda_wait(c.ctx, 12345) < 0                              # unknown ticket
    assert lib.vp9cuda_run_staged(c.ctx, 3, None, None, None) < 0           # unknown handle
    c.close()


def test_device_resident_output_matches_host_copy(workdir):
    """vp9cuda_map_frame: the planes left in HBM are the picture the D2H path delivers."""
    import torch
    from ffmpeg_hybrid_b200 import binding
    import ctypes as C
    lib = binding.load()
    ivf = parity.write_stream(configs.small_variants()["c3_small"], workdir / "map.ivf")
    parity.dump_records(ivf, workdir / "map.v9r")
    frames = binding.load_stream(lib, workdir / "map.v9r", pinned=True)
    ctx = binding.Context(0)
    cudart = C.CDLL("libcudart.so.12")
    cudart.cudaMemcpy2D.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int]
    checked = 0
    for f in frames:
        out = binding.OutFrame(lib, f.w, f.h, f.bpp)
        ctx.wait(ctx.decode(f, out))
        m = ctx.map_frame(-1)
        assert (m.w, m.h, m.bpp) == (f.w, f.h, f.bpp)
        bps = 2 if f.bpp > 8 else 1
        got = b""
        for p in range(3):
            w, h = (f.w + (p > 0)) >> (p > 0), (f.h + (p > 0)) >> (p > 0)
            buf = (C.c_ubyte * (w * h * bps))()
            assert cudart.cudaMemcpy2D(buf, w * bps, m.planes[p], m.strides[p], w * bps, h, 2) == 0
            got += bytes(buf)
        ctx.unmap_frame(m)
        assert got == out.bytes()
        checked += 1
        out.free()
    assert checked == len(frames)
    assert lib.vp9cuda_unmap_frame(ctx.ctx, 1234) < 0          # not a mapped plane
    ctx.close()
    for f in frames:
        f.free()


def test_inter_prediction_behind_previous_loop_filter_agrees(workdir):
    """set_option("overlap_inter", 1): frame N+1's inter prediction is launched in bands behind one-warp
    gates on frame N's loop-filter rows (k_inter_gate); same pictures as the serial order."""
    from ffmpeg_hybrid_b200 import binding
    lib = binding.load()
    for name in ("c3_small", "c2_small", "tilerows"):
        ivf = parity.write_stream(configs.small_variants()[name], workdir / f"ov_{name}.ivf")
        parity.dump_records(ivf, workdir / f"ov_{name}.v9r")
        frames = binding.load_stream(lib, workdir / f"ov_{name}.v9r", pinned=True)
        sums = {}
        for ov, bands in ((0, 6), (1, 6), (1, 1), (1, 64)):
            ctx = binding.Context(0)
            ctx.option("overlap_inter", ov)
            ctx.option("inter_bands", bands)
            handles = [ctx.stage(f) for f in frames]
            got = []
            for _ in range(3):
                tickets = [ctx.run_staged(h) for h in handles]     # back to back: frames really overlap
                ctx.wait(tickets[-1])
                got.append(ctx.checksum(-1))
                for i, h in enumerate(handles):                     # and one by one for every frame's picture
                    ctx.wait(ctx.run_staged(h))
                    got.append(ctx.checksum(-1))
            sums[(ov, bands)] = got
            ctx.close()
        assert len(set(map(tuple, sums.values()))) == 1, name
        for f in frames:
            f.free()


def test_fused_and_separate_wavefront_launches_agree(workdir):
    """k_wave (intra + loop-filter wavefronts in one launch) against separate k_intra / k_lf launches."""
    from ffmpeg_hybrid_b200 import binding
    lib = binding.load()
    ivf = parity.write_stream(configs.small_variants()["c2_small"], workdir / "fw.ivf")
    parity.dump_records(ivf, workdir / "fw.v9r")
    frames = binding.load_stream(lib, workdir / "fw.v9r", pinned=True)
    sums = {}
    for fuse in (1, 0):
        ctx = binding.Context(0)
        ctx.option("fuse_wave", fuse)
        handles = [ctx.stage(f) for f in frames]
        got = []
        for _ in range(3):                       # timing varies from run to run; the pictures must not
            for h in handles:
                ctx.wait(ctx.run_staged(h))
                got.append(ctx.checksum(-1))
        sums[fuse] = got
        ctx.close()
    assert sums[0] == sums[1]
    for f in frames:
        f.free()
