#!/bin/bash
# Development: build tools/build/librfk_<name>.so with extra -D flags on fused_trial.cu (or RFK_VARIANT_SRC=fused_pair) (the other objects
# are reused from rafiki_b200/csrc/build).  Select at run time with RFK_LIB=tools/build/librfk_<name>.so.
set -e
name=$1; shift
src=${RFK_VARIANT_SRC:-fused_trial}
root=$(cd "$(dirname "$0")/.." && pwd)
csrc=$root/rafiki_b200/csrc
out=$root/tools/build
mkdir -p $out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xptxas -v "$@" \
  -c $csrc/$src.cu -o $out/fused_$name.o 2> $out/fused_$name.ptxas.log || (cat $out/fused_$name.ptxas.log; exit 1)
objs=$(ls $csrc/build/*.o | grep -v $src.o)
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/librfk_$name.so $objs $out/fused_$name.o -cudart static -ldl
grep -E "spill|registers" $out/fused_$name.ptxas.log | tail -2
echo built $out/librfk_$name.so
