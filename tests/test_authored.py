"""Syntax-driven streams (ffmpeg_hybrid_b200/author.py, tools/vp9author.c): the reference's own parser run as a
writer, so that motion vectors are authored ON PURPOSE -- final vectors uniform in a chosen range, encoded as
differences against the find_ref_mvs predictor -- instead of sampled.  CPU: the writer is deterministic, the
unmodified reference decodes its output (err_detect explode) to the committed lists with 1 and 4 threads, every
vector lies in the authored range, prediction windows leave the frame on all four sides, the restatement and the
host-emulated CUDA inter kernel reproduce the reference.  GPU: the overlay is bit-exact on them."""
import hashlib
import json
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

from ffmpeg_hybrid_b200 import author, parity, v9r

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tools"))
import record_coverage  # noqa: E402

GOLD = json.loads((ROOT / "tests" / "golden" / "authored.json").read_text())
NAMES = sorted(author.authored_variants())


@pytest.fixture(scope="module")
def streams(workdir):
    out = {}
    for name, (spec, policy) in author.authored_variants().items():
        out[name] = author.author_stream(spec, policy, workdir / f"{name}.ivf")
    return out


@pytest.mark.parametrize("name", NAMES)
def test_writer_is_deterministic_and_reference_decodes(name, streams):
    assert hashlib.sha256(streams[name].read_bytes()).hexdigest() == GOLD[name]["ivf_sha256"]
    assert parity.ref_md5(streams[name]) == GOLD[name]["framemd5"]
    assert parity.ref_md5(streams[name], threads=4) == GOLD[name]["framemd5"]


@pytest.mark.parametrize("name", NAMES)
def test_vectors_are_the_authored_ones(name, streams, workdir):
    rng = int(dict(kv.split("=") for kv in GOLD[name]["policy"].split(","))["mv_range"])
    dump = workdir / f"{name}.v9r"
    parity.dump_records(streams[name], dump)
    mvs = []
    for fr in v9r.read_v9r(dump):
        for lane in fr.lanes:
            b = np.frombuffer(lane.blk, v9r.BLK_DTYPE)
            for k in range(len(b)):
                n = 2 if b["comp"][k] else 1
                mvs.append(b["mv"][k][:, :n, :].reshape(-1, 2))
    mv = np.concatenate(mvs).astype(int)
    assert len(mv) > 500 and np.abs(mv).max() <= rng          # every vector (NEWMV, or inherited from one) is in range
    assert np.abs(mv).max() > rng * 0.9                       # ... and the range is used
    cov = record_coverage.coverage_of([dump])
    assert all(cov.offframe[s] >= 20 for s in ("left", "top", "right", "bottom")), cov.offframe
    if name == "authored_hp":
        assert len({(a, b) for (c, a, b) in cov.phases if c == 0}) == 64
    if name == "authored_scaled":
        assert sum(v for k, v in cov.slots.items() if k.startswith("smc")) > 100


@pytest.mark.parametrize("name", NAMES)
def test_model_and_emulated_kernel_on_authored_streams(name, streams, workdir):
    dump, ref_yuv, mod_yuv = workdir / f"{name}_m.v9r", workdir / f"{name}_ref.yuv", workdir / f"{name}_mod.yuv"
    parity.dump_records(streams[name], dump)
    subprocess.run([str(parity.REF_BIN), "-raw", str(ref_yuv), str(streams[name])], check=True, capture_output=True)
    subprocess.run([str(parity.MODEL_BIN), str(dump), str(mod_yuv)], check=True, capture_output=True)
    assert ref_yuv.read_bytes() == mod_yuv.read_bytes()
    p = subprocess.run([str(ROOT / "oracle" / "_ref" / "emu_inter"), str(dump), "3"], capture_output=True, text=True, timeout=900)
    assert p.returncode == 0 and " 0 differing" in p.stderr, p.stderr[-800:]


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("threads,ttype", [(1, None), (4, "frame")])
def test_overlay_bit_exact_on_authored_streams(name, threads, ttype, streams):
    got = parity.cuda_md5(streams[name], threads=threads, thread_type=ttype)
    assert got == GOLD[name]["framemd5"], parity.compare(GOLD[name]["framemd5"], got)[1]
