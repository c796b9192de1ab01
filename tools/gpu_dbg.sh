#!/bin/bash
# development: check + time + profile the pair kernel
O=gpurun_out/${1:-dbg}; mkdir -p $O; rm -f $O/*
timeout 300 python tools/pair_check.py > $O/pair_check.log 2>&1; echo "pair_check rc=$?"; grep -v "timed out" $O/pair_check.log | tail -4
RFK_ITERS=50 timeout 120 python tools/time_fused.py 8x65536 4x65536 2x65536 1x65536 1x20000 2>&1 | grep -v "timed out" | tee $O/time.log
RFK_PAIR_PROF=$O/prof_w8.txt RFK_ITERS=3 timeout 120 python tools/time_fused.py 8x65536 > /dev/null 2>&1
python tools/prof_pair.py $O/prof_w8.txt | grep "cta0"
RFK_ITERS=1 RFK_PAIR_TRACE=$O/trace_w8.txt timeout 120 python tools/time_fused.py 8x65536 > $O/trace_w8.log 2>&1
python tools/trace_tokens.py $O/trace_w8.txt 104 | head -20
