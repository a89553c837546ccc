"""The CUDA inter-prediction kernel (csrc/k_inter_sb.cuh: reference windows staged in shared memory,
horizontal / vertical 8-tap passes, compound averaging, scaled references, fused inter residual) run
ON THE HOST from the same source through oracle/cuda_emu.h (every CUDA thread an OS thread, real
barriers), compared sample by sample with phases 1-2 of the oracle restatement on every frame of
the parity variants.  The GPU tests then pin the real thing against the reference decoder."""
import subprocess
from pathlib import Path

import pytest

from ffmpeg_hybrid_b200 import configs, parity

ROOT = Path(__file__).resolve().parent.parent
EMU = ROOT / "oracle" / "_ref" / "emu_inter"
BIG = {"manytiles", "wide64", "bilinear", "scale_filters", "c4_small", "c3n_small", "p1_444", "p3_444_10", "p3_444_12"}      # emulated in part (first frames) to keep the suite short
SMALL = sorted(configs.small_variants())


@pytest.mark.parametrize("name", SMALL)
def test_emulated_kernel_equals_restatement(name, workdir):
    ivf = parity.write_stream(configs.small_variants()[name], workdir / f"emu_{name}.ivf")
    v9r = workdir / f"emu_{name}.v9r"
    parity.dump_records(ivf, v9r)
    cmd = [str(EMU), str(v9r), "2" if name in BIG else "3"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-1500:]
    assert " 0 differing" in p.stderr


@pytest.mark.parametrize("name", ["libvpx_cif", "libvpx_odd"])
def test_emulated_kernel_on_encoded_streams(name, workdir):
    v9r = workdir / f"emu_{name}.v9r"
    parity.dump_records(ROOT / "tests" / "golden" / f"{name}.ivf", v9r)
    p = subprocess.run([str(EMU), str(v9r), "6"], capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-1500:]
    assert " 0 differing" in p.stderr


def test_tile_threaded_records(workdir):
    """several producer lanes (one per tile column): lane bases of blocks, residual blocks and coefficients"""
    import os
    ivf = parity.write_stream(configs.small_variants()["c2_small"], workdir / "emu_mt.ivf")
    v9r = workdir / "emu_mt.v9r"
    subprocess.run([str(parity.CUDA_BIN), "-threads", "4", "-thread_type", "slice", "-null", str(ivf)], check=True,
                   capture_output=True, env=dict(os.environ, VP9CUDA_DUMP=str(v9r)))
    p = subprocess.run([str(EMU), str(v9r)], capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-1500:]


@pytest.mark.parametrize("name,frames", [("tiny", 4), ("hbd12", 2), ("lossless", 1), ("intraonly", 2), ("c4_small", 1),
                                         ("p1_444", 1), ("p3_444_12", 2)])
def test_whole_gpu_pipeline_emulated(name, frames, workdir):
    """-full: every kernel of a frame (k_inter_sb, k_intra_prep, k_intra_res, k_lf_sched and the intra /
    loop-filter wavefronts of k_wave, CTAs in block-index order) on the host; the finished picture must equal
    the restatement's.  This is the functional check of the wavefront kernels' indexing and hand-overs that
    needs no GPU; the races between concurrently running CTAs are what the GPU tests add."""
    ivf = parity.write_stream(configs.small_variants()[name], workdir / f"emuf_{name}.ivf")
    v9r = workdir / f"emuf_{name}.v9r"
    parity.dump_records(ivf, v9r)
    p = subprocess.run([str(EMU), "-full", str(v9r), str(frames)], capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-1500:]
    assert " 0 differing" in p.stderr
