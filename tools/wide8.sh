#!/bin/bash
# usage: tools/wide8.sh NGPUS : step time of the widened model (H=4096) with the default bucket plan, 1 and 2 comm streams
N=${1:-8}
mkdir -p gpurun_out
for S in 1 2; do
  DDPB_COMM_STREAMS=$S DDPB_BUCKET_DIAG=0 timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29720+S)) tools/bench_wide.py --hidden 4096 --steps 30 --bucket-caps 0 > gpurun_out/wide8_s$S.jsonl 2> gpurun_out/wide8_s$S.err
  echo "comm streams $S"; python tools/print_bucket_diag.py gpurun_out/wide8_s$S.jsonl
done
