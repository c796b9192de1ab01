#!/bin/bash
# round-2 call 1: determinism stress + timing of issuer variants, then the GPU suite without -x
O=gpurun_out/call1; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
run() { # name lib debug
  echo "=== $1" | tee -a $O/stress.log
  RFK_LIB=$2 RFK_FUSED_DEBUG=$3 timeout 300 python tools/stress_determinism.py --reps 20 --hetero --load >> $O/stress.log 2>&1
  echo "rc=$?" >> $O/stress.log
  RFK_LIB=$2 RFK_FUSED_DEBUG=$3 RFK_ITERS=50 timeout 120 python tools/time_fused.py 1x65536 2x65536 8x65536 3x4096 >> $O/time.log 2>&1
}
D=rafiki_b200/librafiki_b200.so
U=tools/build/librfk_uw.so
run base $D 0
run solo $D 8
run uw $U 0
run uw_solo $U 8
timeout 900 python -m pytest tests -q -m gpu > $O/pytest.log 2>&1
echo "pytest rc=$?" >> $O/pytest.log
tail -3 $O/pytest.log
grep -E "STRESS|===|rc=" $O/stress.log
cat $O/time.log
