#!/bin/bash
# On the GPU box: phase stamps of one mid-horizon general sweep at C4.  Usage: scripts/gpu_sweep_trace.sh <tag>
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
LSMC_TRACE=$OUT/sweep_trace_$TAG.bin python bench.py --config c4 --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-panels > $OUT/sweep_trace_bench_$TAG.json 2> $OUT/sweep_trace_$TAG.err
echo "exit $?"; python scripts/sweep_trace_report.py $OUT/sweep_trace_$TAG.bin | tee $OUT/sweep_trace_$TAG.txt
