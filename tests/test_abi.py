"""The C-ABI library loads and exports every symbol include/vp9cuda.h declares; record structs
have the sizes the header documents; without a GPU the product path fails loudly (-ENODEV)."""
import ctypes as C
import re
from pathlib import Path

import pytest

from ffmpeg_hybrid_b200 import binding, v9r

ROOT = Path(__file__).resolve().parent.parent
HEADER = (ROOT / "include" / "vp9cuda.h").read_text()


def declared_symbols():
    return sorted(set(re.findall(r"\b(vp9cuda_[a-z_0-9]+)\s*\(", HEADER)))


def test_every_declared_symbol_is_exported():
    lib = binding.load()
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"libvp9cuda.so does not export {s}"
    assert sorted(binding.API) == syms


def test_struct_sizes_match_header():
    assert v9r.BLK_SIZE == 48 and v9r.TX_SIZE == 16 and v9r.ITX_SIZE == 20
    assert v9r.SBI_SIZE == 16 and v9r.LF_SIZE == 192
    # header comments state the sizes; the dump writer stores sizeof() and the reader checks them
    for n, sz in (("Vp9CudaInterBlock", 48), ("Vp9CudaTx", 16), ("Vp9CudaIntraTx", 20), ("Vp9CudaSbIndex", 16),
                  ("Vp9CudaLoopFilterSb", 192), ("Vp9CudaCoef16", 4), ("Vp9CudaCoef32", 8)):
        assert re.search(r"}\s*%s;\s*/\*\s*%d bytes" % (n, sz), HEADER), n
    assert C.sizeof(binding.LaneCounts) == v9r.COUNTS_DTYPE.itemsize == 16
    assert C.sizeof(binding.LaneRing) == 72


def test_version_string():
    v = binding.load().vp9cuda_version()
    assert b"sm_100a" in v
    abi = int(re.search(r"#define VP9CUDA_ABI_VERSION\s+(\d+)", HEADER).group(1))
    assert v.endswith(b"abi %d" % abi), v


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = binding.load()
    ctx = C.c_void_p()
    rc = lib.vp9cuda_create(C.byref(ctx), 0, 0, 0, 12, 2)
    assert rc < 0 and not ctx.value
    import numpy as np
    rc = lib.vp9cuda_slot_loop_filter(0, 8, 0, None, None, None, None, None, None, None)
    assert rc < 0


def test_product_does_not_touch_oracle():
    """nothing under ffmpeg_hybrid_b200/csrc or host/ may reference oracle/ (checker only)"""
    for p in list((ROOT / "ffmpeg_hybrid_b200" / "csrc").glob("*")) + list((ROOT / "ffmpeg_hybrid_b200" / "host").glob("*.[ch]")):
        if p.is_file() and p.suffix in (".cu", ".cuh", ".h", ".c"):
            txt = p.read_text()
            assert "replay_model" not in txt or "oracle/" in txt and "#include" not in txt.split("replay_model")[0][-40:], p
            assert '#include "../oracle' not in txt and "dlopen(\"oracle" not in txt, p
