import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); calls through the C ABI")


def _have_reference_tree() -> bool:
    return (Path(os.environ.get("VP9CUDA_REFERENCE", "/root/reference")) / "libavcodec" / "vp9.c").exists()


@pytest.fixture(scope="session", autouse=True)
def built():
    """Make sure every binary exists.  In the build container (reference tree present) missing
    pieces are compiled; on the GPU box they arrive prebuilt with the snapshot."""
    need = [ROOT / "ffmpeg_hybrid_b200" / "libvp9cuda.so", ROOT / "oracle" / "_ref" / "ffvp9_ref",
            ROOT / "oracle" / "_ref" / "libvp9dsp_ref.so", ROOT / "oracle" / "_ref" / "libreplay_model.so",
            ROOT / "oracle" / "_ref" / "replay_model", ROOT / "ffmpeg_hybrid_b200" / "host" / "_build" / "ffvp9_cuda", ROOT / "oracle" / "_ref" / "emu_inter", ROOT / "oracle" / "_ref" / "author" / "ffvp9_author",
            ROOT / "oracle" / "_ref" / "ffvp9_multi_ref", ROOT / "ffmpeg_hybrid_b200" / "host" / "_build" / "ffvp9_multi"]
    if not all(p.exists() for p in need):
        if not _have_reference_tree() and not (ROOT / "oracle" / "_ref" / "ffvp9_ref").exists():
            pytest.exit("oracle/_ref is missing and there is no reference tree to build it from", 2)
        subprocess.run([sys.executable, "-c", "import __graft_entry__ as g; g.build()"], cwd=ROOT, check=True)
    return True


@pytest.fixture(scope="session")
def workdir(tmp_path_factory):
    return tmp_path_factory.mktemp("vp9")
