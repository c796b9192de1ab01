#!/bin/bash
O=gpurun_out/call6; mkdir -p $O
nvidia-smi -L > $O/smi.txt
timeout 600 python -m pytest tests -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log; tail -3 $O/pytest.log
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench1 rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench2 rc=$?"
tail -5 $O/bench_n2.err
cut -c1-1500 $O/bench_n1.json; cut -c1-1800 $O/bench_n2.json
