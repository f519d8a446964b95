#!/bin/bash
# Runs on the GPU box (under gpurun): parity check, bench line, ncu launch list and full captures of the hot kernels.
# Usage: scripts/gpu_profile.sh <tag>
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi_$TAG.csv 2>&1
python bench.py --steps 5 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
echo "bench exit $?"; cat $OUT/bench_$TAG.json
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 1106 -c 1106 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launch_$TAG.log 2>&1
echo "launch list exit $?"
$CMD > $OUT/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"sweep4_kernel|sweep_kernel" -s 500 -c 2 -o $OUT/sweep_$TAG -f $CMD > $OUT/ncu_sweep_$TAG.log 2>&1
echo "sweep capture exit $?"
$CMD > $OUT/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"forward_fast_kernel|forward_kernel|simulate_pair_kernel|simulate_kernel" -s 1 -c 3 -o $OUT/fwdsim_$TAG -f $CMD > $OUT/ncu_fwdsim_$TAG.log 2>&1
echo "fwd/sim capture exit $?"
