#!/bin/bash
# Development: builds variants of one translation unit with extra -D flags into scratch_ab/liblsmc_<name>.so (FAST shapes).
# Usage: scripts/build_variants.sh <unit: kernels|sweep4> name1:"-DFOO=1" name2:"-DBAR=2" ...; run with LSMC_B200_LIB=scratch_ab/liblsmc_<name>.so
set -e
cd "$(dirname "$0")/../storage_b200/csrc"
UNIT=$1; shift
make FAST=1 >/dev/null
mkdir -p ../../scratch_ab
FLAGS="-DLSMC_FAST_BUILD -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -ccbin /usr/bin/g++ -Xcompiler -fPIC,-ffp-contract=off,-Wall -fmad=false"
for spec in "$@"; do
  name=${spec%%:*}; defs=${spec#*:}
  ( OBJS=""; for u in lsmc_api kernels sweep sweep4 host_plan noalloc tail forward2; do if [ $u == $UNIT ]; then OBJS="$OBJS ../../scratch_ab/${UNIT}_$name.o"; else OBJS="$OBJS build_fast/$u.o"; fi; done
    nvcc $FLAGS $defs -c -o ../../scratch_ab/${UNIT}_$name.o $UNIT.cu &&
    nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../scratch_ab/liblsmc_$name.so $OBJS -cudart static -ldl ) &
done
wait
ls ../../scratch_ab/*.so
