#!/bin/bash
# ncu evidence: launch list of a short bench run, full captures of the fused pair kernel and of the memory-bound kernels
O=gpurun_out/${1:-ncu}; mkdir -p $O; rm -f $O/*
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
echo "launch list rc=$?"
$CMD > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fused_pair -s 2 -c 1 -o $O/fused_pair $CMD > $O/ncu_pair.log 2>&1
echo "pair capture rc=$?"
SM="python bench.py --workload small --steps 5 --warmup 1"
$SM > $O/small_plain.log 2>&1 &&
ncu --set full --clock-control none -k regex:'decode_unit|resize_u8_staged|ensemble_mean_vec4' -s 4 -c 12 -o $O/small $SM > $O/ncu_small.log 2>&1
echo "small capture rc=$?"
tail -2 $O/small_plain.log | cut -c1-1500
ls -la $O
