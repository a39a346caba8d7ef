#!/bin/bash
# step time of the widened model on one GPU per DDPB_WIDE_ADAM mode
for mode in end fused; do
  echo "mode=$mode"
  DDPB_WIDE_ADAM=$mode timeout 120 python tools/bench_wide.py --hidden 1024 2048 4096 8192 --steps 20 2>/dev/null | python tools/print_wide.py
done
