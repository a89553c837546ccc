"""Frame-level parity driver: reference software decoder vs the vp9cuda back end.

Both sides are the SAME libavcodec "vp9" decoder front end (tools/ivf_framemd5.c); one is linked
against the unmodified reference objects (oracle/_ref/ffvp9_ref, test infrastructure), the other
against the overlay build whose reconstruction runs in libvp9cuda.so.  The comparison is the
per-frame MD5 list, i.e. `ffmpeg -f framemd5` (reference tests/fate-run.sh:172-174).
"""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path
from typing import List, Optional, Tuple

ROOT = Path(__file__).resolve().parent.parent
REF_BIN = ROOT / "oracle" / "_ref" / "ffvp9_ref"
CUDA_BIN = ROOT / "ffmpeg_hybrid_b200" / "host" / "_build" / "ffvp9_cuda"
CUDA_LIB = ROOT / "ffmpeg_hybrid_b200" / "libvp9cuda.so"
MODEL_BIN = ROOT / "oracle" / "_ref" / "replay_model"


def write_stream(spec, path: Path) -> Path:
    from .streamgen import encode_stream
    path.parent.mkdir(parents=True, exist_ok=True)
    path.write_bytes(encode_stream(spec))
    return path


def _run(cmd: List[str], env=None, timeout=None) -> Tuple[int, str, str]:
    timeout = timeout or float(os.environ.get("VP9CUDA_TEST_TIMEOUT", "600"))
    p = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=timeout)
    return p.returncode, p.stdout, p.stderr


def ref_md5(ivf: Path, threads: int = 1, allow_errors: bool = False) -> List[str]:
    rc, out, err = _run([str(REF_BIN), "-threads", str(threads), str(ivf)])
    if allow_errors:      # damaged streams: the pictures that came out + the packets that failed
        return out.strip().splitlines() + [l for l in err.splitlines() if l.startswith("decode error")]
    if rc:
        raise RuntimeError(f"reference decoder failed on {ivf}: {err[-400:]}")
    return out.strip().splitlines()


def cuda_md5(ivf: Path, device: int = 0, threads: int = 1, thread_type: Optional[str] = None,
             allow_errors: bool = False) -> List[str]:
    env = dict(os.environ, VP9CUDA_LIB=str(CUDA_LIB), VP9CUDA_DEVICE=str(device))
    env.pop("VP9CUDA_DUMP", None)
    cmd = [str(CUDA_BIN), "-threads", str(threads)]
    if thread_type:
        cmd += ["-thread_type", thread_type]
    rc, out, err = _run(cmd + [str(ivf)], env=env)
    if allow_errors:
        return out.strip().splitlines() + [l for l in err.splitlines() if l.startswith("decode error")]
    if rc:
        raise RuntimeError(f"vp9cuda decoder failed on {ivf}: {err[-800:]}")
    return out.strip().splitlines()


def dump_records(ivf: Path, v9r: Path) -> None:
    env = dict(os.environ, VP9CUDA_DUMP=str(v9r))
    rc, out, err = _run([str(CUDA_BIN), "-null", str(ivf)], env=env)
    if rc:
        raise RuntimeError(f"record dump failed on {ivf}: {err[-400:]}")


def compare(a: List[str], b: List[str]) -> Tuple[bool, str]:
    if len(a) != len(b):
        return False, f"frame count {len(a)} vs {len(b)}"
    for i, (x, y) in enumerate(zip(a, b)):
        if x != y:
            return False, f"first mismatch at output frame {i}: {x} vs {y}"
    return True, f"{len(a)} frames identical"


def main(argv: List[str]) -> int:
    from . import configs
    out = Path(argv[1]) if len(argv) > 1 else Path("/tmp/vp9cuda_parity")
    names = argv[2:]
    bad = 0
    specs = dict(configs.small_variants())
    full = {"c1": configs.c1, "c2": configs.c2, "c3": configs.c3, "c4": configs.c4,
            "c5": lambda: configs.c5(0)}
    for n in names:
        if n in full:
            specs[n] = full[n]()
    for n, spec in specs.items():
        if names and n not in names:
            continue
        ivf = write_stream(spec, out / f"{n}.ivf")
        try:
            ok, msg = compare(ref_md5(ivf), cuda_md5(ivf))
        except Exception as e:  # noqa: BLE001
            ok, msg = False, str(e)
        print(f"{n:12s} {'OK  ' if ok else 'FAIL'} {msg}", flush=True)
        bad += not ok
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
