#!/bin/bash
N=${1:-2}
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}" 2>&1 | grep -v "OMP_NUM\|\*\*\*\|^$\|NCCL version"; }
RFK_TRIALS=8 run 29586 tools/time_peer.py
RFK_TRIALS=8 RFK_PEER_NO_MULTI_SCATTER=1 run 29587 tools/time_peer.py | sed 's/^/fixed-root /'
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29588 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('bench N=$N', d['value']/1e6, d['ms_per_step'], d.get('parity_vs_single_gpu'), d['e2e']['value']/1e6, d['gpu_launches'])"
