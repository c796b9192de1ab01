#!/bin/bash
O=gpurun_out/${1:-peer2}; mkdir -p $O
N=${2:-2}
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}" 2>&1 | grep -v "OMP_NUM\|\*\*\*\|^$\|NCCL version"; }
run 29582 tools/check_peer_ensemble.py 20000 | tail -4
run 29583 tools/check_peer_ensemble.py 20000 $N | tail -8
RFK_TRIALS=$N run 29584 tools/time_peer.py | tee -a $O/time_peer_n$N.log
run 29585 tools/check_peer_ensemble.py 20000 3 | tail -3
RFK_TRIALS=8 run 29586 tools/time_peer.py | tee -a $O/time_peer_n${N}_w8.log
