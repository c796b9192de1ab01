#!/bin/bash
O=gpurun_out/${1:-misc}; mkdir -p $O; rm -f $O/*
python tools/sanitize_small.py > $O/sanitize_plain.log 2>&1 && timeout 600 compute-sanitizer --tool memcheck python tools/sanitize_small.py > $O/sanitize_memcheck.log 2>&1
echo "memcheck rc=$?"; tail -4 $O/sanitize_memcheck.log
timeout 600 python bench.py --workload c5 --steps 200 > $O/bench_c5.json 2> $O/bench_c5.err; echo "c5 rc=$?"
timeout 300 python bench.py --workload c2 --steps 200 --warmup 10 > $O/bench_c2.json 2> $O/bench_c2.err; echo "c2 rc=$?"
timeout 300 python bench.py --workload c4 --steps 50 --warmup 5 > $O/bench_c4.json 2> $O/bench_c4.err; echo "c4 rc=$?"
cut -c1-700 $O/bench_c2.json; cut -c1-500 $O/bench_c4.json; cut -c1-400 $O/bench_c5.json
