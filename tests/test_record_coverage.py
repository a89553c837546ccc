"""What the parity streams exercise, counted instead of assumed (tools/record_coverage.py): every
VP9DSPContext slot a 4:2:0 stream can reach (reference libavcodec/vp9dsp.h:40-122 -- the list
checkasm enumerates, tests/checkasm/vp9dsp.c:50-625) must occur at least ten times in the record
streams of the small variants plus the libvpx-encoded streams, every sub-pel phase pair must occur,
and prediction windows must cross all four frame edges."""
import sys
from pathlib import Path

import pytest

from ffmpeg_hybrid_b200 import configs, parity

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tools"))
import record_coverage  # noqa: E402


@pytest.fixture(scope="module")
def cov(workdir):
    paths = []
    for name, spec in configs.small_variants().items():
        if (spec.ss_h, spec.ss_v) != (1, 1):
            continue                          # the slot inventory is the 4:2:0 one (chroma block sizes and vector units)
        ivf = parity.write_stream(spec, workdir / f"cov_{name}.ivf")
        parity.dump_records(ivf, workdir / f"cov_{name}.v9r")
        paths.append(workdir / f"cov_{name}.v9r")
    for ivf in sorted((ROOT / "tests" / "golden").glob("libvpx_*.ivf")):
        parity.dump_records(ivf, workdir / f"cov_{ivf.stem}.v9r")
        paths.append(workdir / f"cov_{ivf.stem}.v9r")
    c = record_coverage.coverage_of(paths)
    print(c.report())
    return c


def test_every_reachable_slot_is_hit(cov):
    missing = cov.missing(10)
    assert not missing, "slots hit fewer than 10 times: " + ", ".join(f"{m} ({cov.slots[m]})" for m in missing)


def test_phases_edges_and_eob_classes(cov):
    assert len({(mx, my) for (c, mx, my) in cov.phases if c == 0}) == 64      # luma: 1/8 pel -> even phases
    assert len({(mx, my) for (c, mx, my) in cov.phases if c == 1}) == 256     # chroma: all 16 x 16
    assert all(cov.offframe[s] >= 100 for s in ("left", "top", "right", "bottom")), cov.offframe
    for slot in record_coverage.all_slots()["itxfm_add"]:
        for cls in ("dc", "sub", "full"):
            assert cov.itx_eob[(slot, cls)] > 0, (slot, cls)
    # every intra mode family meets missing left / top / right edges somewhere
    seen = {(l, t, r) for (_, l, t, r) in cov.intra_edges}
    assert len(seen) == 8
    # scaled prediction in both directions and at a non-integer, anisotropic ratio
    assert {(32, 32), (8, 8), (1, 1)} <= set(cov.steps) and any(dx != dy for dx, dy in cov.steps)


def test_loop_filter_slots_on_both_plane_classes(cov):
    for slot in record_coverage.all_slots()["loop_filter"]:
        assert cov.lf_plane[(slot, "y")] > 0, slot
        if not slot.startswith("loop_filter_16") and "[2]" not in slot.split("]")[0] + "]":
            assert cov.lf_plane[(slot, "uv")] > 0, slot
