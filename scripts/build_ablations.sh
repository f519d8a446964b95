#!/bin/bash
# Development: builds timing-ablation variants of the pipelined sweep into scratch_ab/liblsmc_ab<N>.so (FAST shapes).
# Usage: scripts/build_ablations.sh 0 1 2 ...   then on the GPU box: scripts/run_ablations.sh 0 1 2 ...
set -e
cd "$(dirname "$0")/../storage_b200/csrc"
make FAST=1 >/dev/null
mkdir -p ../../scratch_ab
FLAGS="-DLSMC_FAST_BUILD -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -ccbin /usr/bin/g++ -Xcompiler -fPIC,-ffp-contract=off,-Wall -fmad=false"
for n in "$@"; do
  ( nvcc $FLAGS ${EXTRA:-} -DLSMC_ABLATE=$n -c -o ../../scratch_ab/sweep4_ab$n.o sweep4.cu &&
    nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../scratch_ab/liblsmc_ab$n.so build_fast/lsmc_api.o build_fast/kernels.o build_fast/sweep.o ../../scratch_ab/sweep4_ab$n.o build_fast/host_plan.o -cudart static -ldl ) &
done
wait
ls -la ../../scratch_ab/*.so
