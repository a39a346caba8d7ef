#!/bin/bash
# usage: tools/bucket_diag.sh NGPUS  -> gpurun_out/bucket_diag_N.jsonl (bucket kernels alone, diagnostic modes 0..3)
N=${1:-8}
mkdir -p gpurun_out
CAPS=${2:-1e9}
DDPB_BUCKET_DIAG=${DDPB_BUCKET_DIAG:-0,1,2,3} timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29711 tools/bench_wide.py --hidden 4096 --steps 20 --bucket-caps $CAPS > gpurun_out/bucket_diag_$N.jsonl 2> gpurun_out/bucket_diag_$N.err
python tools/print_bucket_diag.py gpurun_out/bucket_diag_$N.jsonl
