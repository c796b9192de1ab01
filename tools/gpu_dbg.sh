#!/bin/bash
# development: check + time the pair kernel
O=gpurun_out/${1:-dbg}; mkdir -p $O; rm -f $O/*
timeout 300 python tools/pair_check.py > $O/pair_check.log 2>&1; echo "pair_check rc=$?"; grep -v "timed out" $O/pair_check.log | tail -2
RFK_ITERS=50 timeout 120 python tools/time_fused.py 8x65536 4x65536 2x65536 1x65536 1x20000 2>&1 | grep -v "timed out" | tee $O/time.log
