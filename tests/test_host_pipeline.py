"""Host side of the frame pipeline without a GPU: the libavcodec overlay writes its record stream to a
file ($VP9CUDA_DUMP, same C-ABI calls as the back end).  Frame threading (one worker per frame, only
the entropy decode on the host; reference pthread_frame.c, vp9.c:1731-1732, 1821-1824) and tile
threading must produce byte-identical records to the single-threaded decoder, and the CPU
restatement must turn them into the reference's pictures."""
import os
import subprocess
from pathlib import Path

import pytest

from ffmpeg_hybrid_b200 import configs, parity, v9r

ROOT = Path(__file__).resolve().parent.parent


def dump(ivf, out, threads=1, ttype=None, allow_fail=False):
    cmd = [str(parity.CUDA_BIN), "-null", "-threads", str(threads)] + (["-thread_type", ttype] if ttype else [])
    p = subprocess.run(cmd + [str(ivf)], env=dict(os.environ, VP9CUDA_DUMP=str(out)), capture_output=True, text=True)
    assert allow_fail or p.returncode == 0, p.stderr[-300:]
    return out


def _canon(path):
    """everything in a record dump except VP9Filter.level entries, which libavcodec leaves stale for
    unfiltered blocks (vp9block.c:1439-1452: no mask bit points at them, they are never read)"""
    import numpy as np
    out = []
    for f in v9r.read_v9r(path):
        lf = np.frombuffer(f.lf, np.uint8).reshape(-1, 192)[:, 64:].tobytes()
        out.append((f.hdr.tobytes(), [(l.counts.tobytes(), l.coef_base, l.blk, l.tx, l.itx, l.coef) for l in f.lanes],
                    f.sb_index, lf))
    return out


@pytest.mark.parametrize("name", ["c2_small", "c3_small", "c3p_small", "parallel", "superframe", "segment", "keyresize",
                                  "intraonly", "scale_mix"])
def test_frame_threaded_records_equal_single_threaded(name, workdir):
    ivf = parity.write_stream(configs.small_variants()[name], workdir / f"{name}_h.ivf")
    one = _canon(dump(ivf, workdir / f"{name}_1.v9r"))
    for threads in (2, 5, 9):
        assert _canon(dump(ivf, workdir / f"{name}_f{threads}.v9r", threads, "frame")) == one, (name, threads)


@pytest.mark.parametrize("name", ["c2_small", "manytiles", "c3p_small"])
def test_tile_threaded_records_hold_the_same_blocks(name, workdir):
    """one producer lane per tile column: same records, only dealt to more lanes"""
    ivf = parity.write_stream(configs.small_variants()[name], workdir / f"{name}_t.ivf")
    a = v9r.read_v9r(dump(ivf, workdir / f"{name}_t1.v9r"))
    b = v9r.read_v9r(dump(ivf, workdir / f"{name}_t8.v9r", 8, "slice"))
    assert len(a) == len(b)
    for fa, fb in zip(a, b):
        assert len(fb.lanes) >= len(fa.lanes) == 1
        for key in ("blk", "tx", "itx", "coef"):
            assert sum(int(l.counts[key]) for l in fb.lanes) == int(fa.lanes[0].counts[key])
        assert fa.lf == fb.lf


def test_damaged_frames_give_their_ring_set_back(workdir):
    """four broken packets in one stream (vp9.c:1340-1346, 1389-1391): the frames around them are
    recorded exactly as a decoder sees them that never met the broken ones, under every thread model
    (a frame that is begun and not finished must be aborted, or the frames behind it wait for ever)"""
    spec = configs.damaged_variants()["bad_many"]
    ivf = parity.write_stream(spec, workdir / "bad_many_h.ivf")
    one = v9r.read_v9r(dump(ivf, workdir / "bm1.v9r", allow_fail=True))
    n_ok = sum(f.damage is None for f in spec.frames)
    assert len(one) == n_ok
    for threads, ttype in ((4, "slice"), (3, "frame"), (8, "frame")):
        got = v9r.read_v9r(dump(ivf, workdir / f"bm_{ttype}{threads}.v9r", threads, ttype, allow_fail=True))
        assert len(got) == n_ok
        if ttype == "frame":
            assert _canon(workdir / f"bm_{ttype}{threads}.v9r") == _canon(workdir / "bm1.v9r")
