#!/bin/bash
O=gpurun_out/call3; mkdir -p $O
for d in 0 1 2 32 33 35 4 64 65 66 67 96 99 103; do
  RFK_FUSED_DEBUG=$d RFK_ITERS=30 timeout 60 python tools/time_fused.py 8x65536 1x65536 >> $O/switch.log 2>&1
done
cat $O/switch.log
