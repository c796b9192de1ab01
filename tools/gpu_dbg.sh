#!/bin/bash
# development: timing experiments on the pair kernel (RFK_PAIR_DEBUG switches; results of those runs are not valid)
O=gpurun_out/${1:-dbg}; mkdir -p $O; rm -f $O/*
RFK_PAIR_PROF=$O/prof_w8.txt RFK_ITERS=3 timeout 120 python tools/time_fused.py 8x65536 2>&1 | tee -a $O/time.log
python tools/prof_pair.py $O/prof_w8.txt
