"""bench.py's reference arm (the driver runs `bench.py --impl reference`): runs on the host cores
only, prints ONE JSON line with the contract's keys, and reports the faster of libavcodec's two
thread models.  (The vp9cuda arm needs a GPU and is exercised by the driver itself.)"""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_reference_arm_prints_the_contract_line():
    p = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--workload", "c1",
                        "--frames", "4", "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-400:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "decoded_frames_per_sec" and d["unit"] == "frames/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["n_gpus"] == 1
    assert d["e2e"] == {"value": d["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"]
    assert set(d["config"]["fps_by_thread_type"]) == {"frame", "slice"}
    assert d["config"]["thread_type"] in ("frame", "slice")


def test_vp9cuda_arm_refuses_to_run_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        return
    p = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--steps", "1"], capture_output=True, text=True,
                       timeout=300)
    assert p.returncode != 0 and "no CPU path" in (p.stderr + p.stdout)
