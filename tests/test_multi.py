"""One process, many streams / closed GOPs (tools/vp9multi.c).  CPU: the scheduler
linked against the unmodified reference decoder (oracle/_ref/ffvp9_multi_ref) reproduces the single-decoder
framemd5 lists, and its GOP split is the one ffmpeg_hybrid_b200.shard.split_closed_gops computes.  GPU: the same
tool over the overlay + libvp9cuda.so (ffvp9_multi), several decoders in one process."""
import json
import os
import re
import subprocess
from pathlib import Path

import pytest

from ffmpeg_hybrid_b200 import configs, parity, shard

ROOT = Path(__file__).resolve().parent.parent
GOLDEN = json.loads((ROOT / "tests" / "golden" / "framemd5.json").read_text())
MULTI_REF = ROOT / "oracle" / "_ref" / "ffvp9_multi_ref"
MULTI_CUDA = ROOT / "ffmpeg_hybrid_b200" / "host" / "_build" / "ffvp9_multi"


def _packets(ivf: Path):
    data = ivf.read_bytes()
    pos, out = 32, []
    while pos + 12 <= len(data):
        n = int.from_bytes(data[pos:pos + 4], "little")
        out.append(data[pos + 12:pos + 12 + n])
        pos += 12 + n
    return out


def _is_key(pkt: bytes) -> bool:
    """frame_marker, profile, show_existing_frame, frame_type (vp9.c:527-552)"""
    bits = "".join(f"{b:08b}" for b in pkt[:2])
    profile = int(bits[2]) | int(bits[3]) << 1
    k = 4 + (1 if profile == 3 else 0)
    return bits[:2] == "10" and bits[k] == "0" and bits[k + 1] == "0"


def _run(binary, args, env=None):
    p = subprocess.run([str(binary)] + args, capture_output=True, text=True, env=env, timeout=600)
    assert p.returncode == 0, p.stderr[-800:]
    return p.stdout.strip().splitlines(), p.stderr


@pytest.mark.parametrize("name,world", [("keyresize", 2), ("c4_small", 2), ("keyresize", 8)])
def test_closed_gops_on_reference_decoder(name, world, workdir):
    ivf = parity.write_stream(configs.small_variants()[name], workdir / f"multi_{name}.ivf")
    flags = [_is_key(p) for p in _packets(ivf)]
    plan = shard.split_closed_gops(flags, world)
    n_gops = sum(len(r) for r in plan)
    assert n_gops == sum(flags) and sorted(g for r in plan for g in r)[0][0] == 0
    out, err = _run(MULTI_REF, ["-gops", "-gpus", str(world), str(ivf)])
    assert out == GOLDEN[name]["framemd5"]
    m = re.search(r"jobs=(\d+) workers=(\d+)", err)
    assert int(m.group(1)) == n_gops and int(m.group(2)) == min(world, n_gops)


def test_streams_side_by_side_on_reference_decoder(workdir):
    names = ["c1_small", "hbd12", "keyresize", "tiny", "odd"]
    ivfs = [parity.write_stream(configs.small_variants()[n], workdir / f"multi_s_{n}.ivf") for n in names]
    out, _ = _run(MULTI_REF, ["-gpus", "2", "-threads", "2"] + [str(p) for p in ivfs])
    got, cur = {}, None
    for line in out:
        if line.startswith("== "):
            cur = line[3:]
            got[cur] = []
        else:
            got[cur].append(line)
    for n, p in zip(names, ivfs):
        assert got[str(p)] == GOLDEN[n]["framemd5"]
    assert shard.streams_of_rank(len(names), 2, 0) == [0, 2, 4]


@pytest.mark.gpu
@pytest.mark.parametrize("name,threads,ttype", [("keyresize", 1, "slice"), ("c4_small", 4, "slice"), ("keyresize", 4, "frame")])
def test_closed_gops_on_gpu(name, threads, ttype, workdir):
    ivf = parity.write_stream(configs.small_variants()[name], workdir / f"multi_g_{name}.ivf")
    env = dict(os.environ, VP9CUDA_LIB=str(parity.CUDA_LIB))
    env.pop("VP9CUDA_DUMP", None)
    out, _ = _run(MULTI_CUDA, ["-gops", "-gpus", "1", "-workers", "2", "-threads", str(threads), "-thread_type", ttype, str(ivf)], env)
    assert out == GOLDEN[name]["framemd5"]


@pytest.mark.gpu
def test_streams_side_by_side_on_gpu(workdir):
    """five decoders with their own back-end contexts in one process on one GPU"""
    names = ["c2_small", "hbd12", "c3_small", "tiny", "odd"]
    ivfs = [parity.write_stream(configs.small_variants()[n], workdir / f"multi_gs_{n}.ivf") for n in names]
    env = dict(os.environ, VP9CUDA_LIB=str(parity.CUDA_LIB))
    env.pop("VP9CUDA_DUMP", None)
    out, _ = _run(MULTI_CUDA, ["-gpus", "1", "-threads", "2", "-thread_type", "frame"] + [str(p) for p in ivfs], env)
    got, cur = {}, None
    for line in out:
        if line.startswith("== "):
            cur = line[3:]
            got[cur] = []
        else:
            got[cur].append(line)
    for n, p in zip(names, ivfs):
        assert got[str(p)] == GOLDEN[n]["framemd5"], n
