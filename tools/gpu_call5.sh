#!/bin/bash
O=gpurun_out/call5; mkdir -p $O
timeout 900 python -m pytest tests -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -3 $O/pytest.log
timeout 300 python tools/stress_determinism.py --reps 25 --hetero --load > $O/stress.log 2>&1; echo "rc=$?" >> $O/stress.log
grep -E "STRESS|rc=" $O/stress.log
RFK_ITERS=50 timeout 120 python tools/time_fused.py 1x65536 2x65536 4x65536 8x65536 3x4096 > $O/time.log 2>&1; cat $O/time.log
RFK_ITERS=1 RFK_FUSED_PROF=2 RFK_FUSED_TRACE_FILE=$O/trace_w8.txt timeout 120 python tools/time_fused.py 8x65536 > $O/prof_w8.log 2>&1
RFK_ITERS=1 RFK_FUSED_PROF=2 RFK_FUSED_TRACE_FILE=$O/trace_w1.txt timeout 120 python tools/time_fused.py 1x65536 > $O/prof_w1.log 2>&1
tail -22 $O/prof_w8.log
