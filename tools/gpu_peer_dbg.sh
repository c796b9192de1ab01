#!/bin/bash
# development: where the pipelined peer step's time goes (RFK_PEER_DEBUG timing switches), N ranks
N=${1:-2}
run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}" 2>&1 | grep -v "OMP_NUM\|\*\*\*\|^$\|NCCL version"; }
port=29600
for d in 0 16 4 20 11 27 31; do
  port=$((port+1))
  RFK_TRIALS=$N RFK_PEER_DEBUG=$d run $port tools/time_peer.py | tail -1
done
RFK_NO_PDL=1 RFK_TRIALS=$N RFK_PEER_DEBUG=0 run 29620 tools/time_peer.py | tail -1 | sed 's/^/nopdl /'
