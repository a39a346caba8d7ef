#!/bin/bash
# One GPU-box visit: smoke, GPU tests, bench, then (only if all of that exited 0) ncu.
# Usage (from the repo root, under gpurun): bash tools/gpu_check.sh [smoke|tests|bench|wide|ncu ...]
set -u
mkdir -p gpurun_out
what="${*:-smoke tests bench wide ncu}"
rc=0
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.csv 2>&1
for step in $what; do
  case $step in
    smoke) python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1 || rc=1; tail -2 gpurun_out/smoke.log ;;
    tests) python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1 || rc=1; tail -15 gpurun_out/pytest_gpu.log ;;
    bench) python bench.py --steps 2000 --warmup 20 > gpurun_out/bench.json 2> gpurun_out/bench.err || rc=1; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
           python bench.py --impl reference --steps 300 --warmup 10 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err || rc=1; cat gpurun_out/bench_ref.json ;;
    wide)  python tools/bench_wide.py > gpurun_out/wide_bench.jsonl 2> gpurun_out/wide_bench.err || rc=1
           cut -c1-330 gpurun_out/wide_bench.jsonl
           [ $rc -eq 0 ] && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gemm_tf32|avg_adam|wide_" -c 120 --csv \
               --log-file gpurun_out/wide_launches.csv python tools/bench_wide.py --hidden 4096 --steps 3 > gpurun_out/ncu_wide.log 2>&1 ;;
    ncu)   [ $rc -eq 0 ] || { echo "skipping ncu: an earlier step failed"; continue; }
           python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
           ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
               python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
           python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 &&
           ncu --set full --clock-control none --import-source on -k regex:toy_step -s 10 -c 3 -f -o gpurun_out/toy_step \
               python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
           tail -3 gpurun_out/ncu_full.log ;;
  esac
done
exit $rc
