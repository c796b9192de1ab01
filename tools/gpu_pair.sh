#!/bin/bash
# development: check, time and trace the CTA-pair kernel; then the GPU suite and the determinism stress
O=gpurun_out/${1:-pair}; mkdir -p $O; rm -f $O/*
timeout 300 python tools/pair_check.py > $O/pair_check.log 2>&1; echo "pair_check rc=$?"; grep -v "timed out" $O/pair_check.log | tail -10; grep -c "timed out" $O/pair_check.log
RFK_ITERS=50 timeout 120 python tools/time_fused.py 8x65536 4x65536 2x65536 1x65536 1x20000 3x4096 2>&1 | grep -v "timed out" | tee $O/time.log
RFK_NO_PAIR=1 RFK_ITERS=50 timeout 120 python tools/time_fused.py 8x65536 1x65536 2>&1 | sed 's/^/nopair /' | tee -a $O/time.log
RFK_PAIR_DEBUG=4 RFK_ITERS=30 timeout 120 python tools/time_fused.py 8x65536 2>&1 | sed 's/^/pairdebug=4 /' | tee -a $O/time.log
if [ "$2" != "quick" ]; then
timeout 900 python -m pytest tests -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log; tail -3 $O/pytest.log
timeout 300 python tools/stress_determinism.py --reps 25 --hetero --load > $O/stress.log 2>&1; echo "rc=$?" >> $O/stress.log; grep -E "STRESS|rc=" $O/stress.log
fi
RFK_ITERS=1 RFK_PAIR_TRACE=$O/trace_w8.txt timeout 120 python tools/time_fused.py 8x65536 > $O/trace_w8.log 2>&1
python tools/trace_pair.py $O/trace_w8.txt > $O/trace_w8_summary.txt 2>&1; head -24 $O/trace_w8_summary.txt
