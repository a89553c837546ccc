#!/bin/bash
# One GPU-box round trip: GPU tests, bench line, ncu launch list and full-set capture.
# usage (through gpurun): bash tools/gpu_round.sh <tag> [pytest-args...]
tag=${1:-x}; shift
cd ${GRAFT_REPO_ROOT:-.}
mkdir -p gpurun_out
nvidia-smi -L | head -1; nproc
timeout 1800 python -m pytest tests -m gpu -x -q "$@" > gpurun_out/r02_pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r02_pytest_gpu_$tag.log
timeout 900 python bench.py > gpurun_out/r02_bench_c3_$tag.json 2> gpurun_out/r02_bench_c3_$tag.err; echo "bench rc=$?"
tail -3 gpurun_out/r02_bench_c3_$tag.err
python - <<PY
import json
d=json.load(open("gpurun_out/r02_bench_c3_$tag.json"))
print("value",d["value"],"e2e",d["e2e"]["value"],"full",d.get("full_decode",{}).get("value"),"cpu",d.get("cpu_baseline",{}).get("value"))
print({k:round(v,3) for k,v in d["roofline"]["stage_ms_per_step"].items()})
print(d.get("concurrent_streams"))
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_ncu_launches_c3_$tag.csv python bench.py --steps 1 --warmup 3 --frames 6 --no-cpu-baseline --no-concurrent --no-full-decode > gpurun_out/ncu_l.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_wave|k_inter|k_itx' -c 14 -o gpurun_out/r02_full_c3_$tag python tools/ncu_frames.py c3 3 > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | tail -12
