#!/bin/bash
# the driver's scaling sequence: N = 1, 2, 4, 8 back to back, 20 steps, 5 warm-up
O=gpurun_out/${1:-scale}; mkdir -p $O; rm -f $O/*
nvidia-smi -L | wc -l
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_n1.json 2> $O/bench_n1.err; echo "n1 rc=$?"
for N in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29550+N)) bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "n$N rc=$?"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --steps 2000 --warmup 20 > $O/bench_n8_long.json 2> $O/bench_n8_long.err; echo "n8 long rc=$?"
python - <<PY
import json
for f in ('bench_n1','bench_n2','bench_n4','bench_n8','bench_n8_long'):
    try:
        for line in open('$O/%s.json' % f):
            if line.startswith('{'):
                d=json.loads(line)
                print(f, 'value %.1f M  ms/step %.4f median %.4f  parity %s  e2e %.1f M (h2d %s)  launches %s' % (d['value']/1e6, d['ms_per_step'], d['ms_median_step'], d.get('parity_vs_single_gpu'), d['e2e']['value']/1e6, d['e2e']['h2d_bytes_per_step'], d['gpu_launches']))
    except Exception as e:
        print(f, 'ERR', e)
PY
tail -3 $O/bench_n8.err
