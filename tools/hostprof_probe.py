import sys, subprocess, os, tempfile
from pathlib import Path
sys.path.insert(0, '.')
import bench
from ffmpeg_hybrid_b200 import parity
tmp = Path(tempfile.mkdtemp())
for wl in ("c3", "c3n"):
    ivf = parity.write_stream(bench.make_spec(wl, 0, 0), tmp / f"{wl}.ivf")
    for label, env in (("gpu", {"VP9CUDA_HOSTPROF": "1"}), ("sink", {"VP9CUDA_DUMP": "/dev/null"}), ("gpu_nopipe", {"VP9CUDA_HOSTPROF": "1", "VP9CUDA_PIPELINE": "0"}),
                       ("gpu_pageable", {"VP9CUDA_HOSTPROF": "1", "VP9CUDA_PINNED_FRAMES": "0"})):
        for rep in range(2):
            p = subprocess.run([str(parity.CUDA_BIN), "-threads", "8", "-thread_type", "slice", "-null", "-loop", "4", str(ivf)],
                               capture_output=True, text=True, env=dict(os.environ, VP9CUDA_LIB=str(parity.CUDA_LIB), **env))
            print(wl, label, rep, p.stderr.strip().replace("\n", " | ")[-600:], flush=True)
