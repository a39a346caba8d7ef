#!/bin/bash
# One visit to an N-GPU box (gpurun --gpus N): multi-GPU tests, the scaling bench at 2..N ranks and
# the bucket-cap sweep of the widened model.  Usage: bash tools/gpu_check8.sh [tests|bench|sweep ...]
set -u
mkdir -p gpurun_out
what="${*:-tests bench sweep}"
ngpu=$(nvidia-smi -L | wc -l)
rc=0
run_n() {  # run_n <ranks> <port> <script and args...>
  local n=$1 port=$2; shift 2
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"
}
for step in $what; do
  case $step in
    tests) timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/pytest_multi.log 2>&1 || rc=1
           tail -4 gpurun_out/pytest_multi.log ;;
    bench) for n in 2 4 8; do
             [ "$n" -le "$ngpu" ] || continue
             run_n $n $((29500 + n)) bench.py --gpus $n --steps 2000 --warmup 20 \
                 > gpurun_out/bench_${n}gpu.json 2> gpurun_out/bench_${n}gpu.err || rc=1
             cat gpurun_out/bench_${n}gpu.json
             DDPB_ALLREDUCE=nccl run_n $n $((29600 + n)) bench.py --gpus $n --steps 500 --warmup 20 \
                 > gpurun_out/bench_${n}gpu_nccl.json 2>> gpurun_out/bench_${n}gpu.err || rc=1
             cut -c1-260 gpurun_out/bench_${n}gpu_nccl.json
           done ;;
    sweep) for n in 2 8; do
             [ "$n" -le "$ngpu" ] || continue
             run_n $n $((29700 + n)) tools/bench_wide.py --hidden 1024 4096 --bucket-caps 1 5 25 100 1000000 \
                 > gpurun_out/wide_sweep_${n}gpu.jsonl 2> gpurun_out/wide_sweep_${n}gpu.err || rc=1
             cat gpurun_out/wide_sweep_${n}gpu.jsonl
           done ;;
  esac
done
exit $rc
