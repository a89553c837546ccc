#!/bin/bash
# Final evidence run of a round on one B200 box: every GPU test, the bench line of every workload, ncu launch list and
# full-set capture of C3, side-by-side streams.   usage (through gpurun): bash tools/gpu_final.sh <tag>
tag=${1:-final}
cd ${GRAFT_REPO_ROOT:-.}
mkdir -p gpurun_out
nvidia-smi -L | head -1; nproc; lscpu | grep "Model name" | head -1
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_gpu_$tag.log
timeout 900 python bench.py > gpurun_out/r02_bench_c3_$tag.json 2> gpurun_out/r02_bench_c3_$tag.err; echo "bench c3 rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/r02_bench_c3_${tag}_reference_arm.json 2>/dev/null; echo "reference arm rc=$?"
for wl in c1 c2 c4 c5 c3n; do
  timeout 900 python bench.py --workload $wl > gpurun_out/r02_bench_${wl}_$tag.json 2> gpurun_out/r02_bench_${wl}_$tag.err; echo "bench $wl rc=$?"
done
python - <<PY
import json
for wl in ("c1","c2","c3","c4","c5","c3n"):
    try:
        d=json.load(open(f"gpurun_out/r02_bench_{wl}_$tag.json"))
        print(wl, "value %.1f e2e %.1f full %.1f ceiling %.1f cpu %.1f conc %s" % (d["value"], d["e2e"]["value"], d.get("full_decode",{}).get("value",0), d.get("entropy_only",{}).get("value",0), d.get("cpu_baseline",{}).get("value",0), {k:round(v) for k,v in d.get("concurrent_streams",{}).items() if k.endswith("streams")}))
    except Exception as e: print(wl, "failed", e)
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_ncu_launches_c3_$tag.csv python bench.py --steps 1 --warmup 3 --frames 6 --no-cpu-baseline --no-concurrent --no-full-decode > gpurun_out/ncu_l.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_wave|k_inter|k_itx' -c 14 -o gpurun_out/r02_full_c3_$tag python tools/ncu_frames.py c3 3 > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
for wl in c3 c3n c2; do python tools/multi_stream.py --workload $wl --streams 1,2,4 --steps 3 2>&1 | tail -1; done > gpurun_out/r02_multi_stream_$tag.log; cat gpurun_out/r02_multi_stream_$tag.log
ls -la gpurun_out | tail -20
