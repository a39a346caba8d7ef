#!/bin/bash
# usage: tools/xchg_timeline.sh NGPUS -> gpurun_out/xchg_timeline_N.jsonl
N=${1:-2}
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741 \
    tools/toy_exchange_timeline.py > gpurun_out/xchg_timeline_$N.jsonl 2> gpurun_out/xchg_timeline_$N.err
echo "rc=$?"; cat gpurun_out/xchg_timeline_$N.jsonl; tail -3 gpurun_out/xchg_timeline_$N.err
