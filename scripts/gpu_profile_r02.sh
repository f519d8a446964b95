#!/bin/bash
# Runs on the GPU box (under gpurun): GPU tests, bench lines (C3 default, C4, C5), ncu launch list and full captures of the hot kernels.
# Usage: scripts/gpu_profile_r02.sh <tag> [notest]
set -u
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi_$TAG.csv 2>&1
if [ "${2:-}" != "notest" ]; then
  python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_gpu_$TAG.log
fi
python bench.py --steps 5 --warmup 3 > $OUT/bench_c3_$TAG.json 2> $OUT/bench_c3_$TAG.err
echo "bench c3 exit $?"; cat $OUT/bench_c3_$TAG.json
python bench.py --config c4 --steps 2 --warmup 1 --no-cpu-baseline --no-panels > $OUT/bench_c4_$TAG.json 2> $OUT/bench_c4_$TAG.err
echo "bench c4 exit $?"; cat $OUT/bench_c4_$TAG.json
python bench.py --config c5 --steps 3 --warmup 2 --no-cpu-baseline --no-panels > $OUT/bench_c5_$TAG.json 2> $OUT/bench_c5_$TAG.err
echo "bench c5 exit $?"; cat $OUT/bench_c5_$TAG.json
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-panels"
$CMD > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launch_$TAG.log 2>&1
echo "launch list exit $?"
ncu --set full --clock-control none --import-source on -k regex:"sweep4_kernel" -s 200 -c 1 -o $OUT/sweep4_$TAG -f $CMD > $OUT/ncu_sweep_$TAG.log 2>&1
echo "sweep capture exit $?"
ncu --set full --clock-control none --import-source on -k regex:"forward2_kernel|simulate_pair_kernel" -c 3 -o $OUT/fwdsim_$TAG -f $CMD > $OUT/ncu_fwdsim_$TAG.log 2>&1
echo "fwd/sim capture exit $?"
CMD4="python bench.py --config c4 --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-panels --paths-total 131072"
$CMD4 > $OUT/plain_c4_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"sweep_kernel" -s 300 -c 1 -o $OUT/sweepg_c4_$TAG -f $CMD4 > $OUT/ncu_sweepg_$TAG.log 2>&1
echo "c4 sweep capture exit $?"
for f in sweep4 fwdsim sweepg_c4; do
  ncu -i $OUT/${f}_$TAG.ncu-rep --page raw --csv > $OUT/${f}_${TAG}_raw.csv 2>/dev/null
  ncu -i $OUT/${f}_$TAG.ncu-rep --page source --csv --print-source sass > $OUT/${f}_${TAG}_source.csv 2>/dev/null
done
du -sh $OUT/*.ncu-rep
find $OUT -name "*.ncu-rep" -size +14M -delete
ls -la $OUT | tail -30
