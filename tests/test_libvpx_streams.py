"""Independently encoded streams (libvpx through OpenCV, tests/golden/make_libvpx_streams.py): what a
real encoder's motion search, mode decision and rate control produce, as a check that does not share
the random-payload generator's blind spots.  CPU: the committed IVF files still decode to the
committed framemd5 lists with the unmodified reference decoder, and the oracle restatement fed
with the glue's records reproduces the reference's pictures.  GPU: the overlay + libvpx9cuda.so
gives the same lists, single-threaded, tile-threaded and frame-threaded."""
import hashlib
import json
import subprocess
from pathlib import Path

import pytest

from ffmpeg_hybrid_b200 import parity

ROOT = Path(__file__).resolve().parent.parent
GOLD = json.loads((ROOT / "tests" / "golden" / "libvpx.json").read_text())
NAMES = sorted(GOLD)


def _ivf(name):
    return ROOT / "tests" / "golden" / f"{name}.ivf"


@pytest.mark.parametrize("name", NAMES)
def test_reference_matches_committed_list(name):
    assert hashlib.sha256(_ivf(name).read_bytes()).hexdigest() == GOLD[name]["ivf_sha256"]
    assert parity.ref_md5(_ivf(name)) == GOLD[name]["framemd5"]


@pytest.mark.parametrize("name", NAMES)
def test_model_equals_reference_on_encoded_streams(name, workdir):
    v9r, ref_yuv, mod_yuv = workdir / f"{name}.v9r", workdir / f"{name}_ref.yuv", workdir / f"{name}_mod.yuv"
    subprocess.run([str(parity.REF_BIN), "-raw", str(ref_yuv), str(_ivf(name))], check=True, capture_output=True)
    parity.dump_records(_ivf(name), v9r)
    subprocess.run([str(parity.MODEL_BIN), str(v9r), str(mod_yuv)], check=True, capture_output=True)
    a, b = ref_yuv.read_bytes(), mod_yuv.read_bytes()
    assert len(a) == len(b) and len(a) > 0
    assert a == b, f"restatement differs from the reference decoder on {name}"


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("threads,ttype", [(1, None), (4, "slice"), (4, "frame")])
def test_overlay_bit_exact_on_encoded_streams(name, threads, ttype):
    got = parity.cuda_md5(_ivf(name), threads=threads, thread_type=ttype)
    assert got == GOLD[name]["framemd5"], parity.compare(GOLD[name]["framemd5"], got)[1]
