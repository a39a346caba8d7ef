#!/bin/bash
# usage: tools/bench_n.sh NGPUS [STEPS] -> gpurun_out/bench_Ngpu_final.json (the driver's launch line)
N=${1:-2}; K=${2:-500}
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29751 \
    bench.py --gpus $N --steps $K --warmup 10 > gpurun_out/bench_${N}gpu_final.json 2> gpurun_out/bench_${N}gpu_final.err
echo "bench rc=$?"
python tools/print_bench.py gpurun_out/bench_${N}gpu_final.json
tail -2 gpurun_out/bench_${N}gpu_final.err
