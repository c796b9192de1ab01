#!/bin/bash
N=${1:-2}
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}" 2>&1 | grep -v "OMP_NUM\|\*\*\*\|^$\|NCCL version"; }
run 29583 tools/check_peer_ensemble.py 20000 $N | tail -2
RFK_TRIALS=$N run 29611 tools/time_peer.py | grep "world"
RFK_PEER_DEBUG=16 RFK_TRIALS=$N run 29612 tools/time_peer.py | grep "RFK_PEER_DEBUG="
