#!/bin/bash
# On the GPU box: the closing evidence of round 2 — GPU test suite, bench lines (C3 with CPU baseline and panels, C3 at the 8-GPU share,
# C4, C5), ncu launch list of a C3 step, full ncu capture of one mid-horizon general sweep and tail kernel at C4, phase stamps of the
# general sweep.  Usage: scripts/gpu_final_r02.sh <tag>
set -u
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_gpu_$TAG.log
python bench.py --steps 10 --warmup 3 > $OUT/bench_c3_$TAG.json 2> $OUT/bench_c3_$TAG.err; echo "bench c3 exit $?"
python bench.py --config c3 --paths-total 125000 --steps 5 --warmup 3 --no-cpu-baseline --no-panels > $OUT/bench_c3_125k_$TAG.json 2> $OUT/bench_c3_125k_$TAG.err
LSMC_TAIL_TRACE=1 python bench.py --config c4 --steps 3 --warmup 1 --no-cpu-baseline --no-panels > $OUT/bench_c4_$TAG.json 2> $OUT/bench_c4_$TAG.err; echo "bench c4 exit $?"
python bench.py --config c5 --steps 3 --warmup 2 --no-cpu-baseline --no-panels > $OUT/bench_c5_$TAG.json 2> $OUT/bench_c5_$TAG.err; echo "bench c5 exit $?"
for n in c3 c3_125k c4 c5; do python -c "
import json;d=json.load(open('$OUT/bench_${n}_$TAG.json'));print('$n', 'ms/step',round(d['ms_per_step'],3),{k:round(v,3) for k,v in d['phase_ms'].items()},'sweep avg ms',round(d['roofline']['avg_launch_ms'],4),'frac',round(d['roofline']['frac'],4),'overhead us',round(d['backward_overhead_us_per_period'],2),'e2e ms',round(d['e2e']['ms_per_step'],2),'npv',d['npv'],'parity',d.get('parity',{}).get('pass'))"; done
grep "tail trace" $OUT/bench_c4_$TAG.err | head -1
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-panels"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launch_$TAG.log 2>&1; echo "launch list exit $?"
CMD4="python bench.py --config c4 --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-panels --paths-total 131072"
ncu --set full --clock-control none --import-source on -k regex:"sweep_kernel|tail_kernel" -s 600 -c 2 -o $OUT/sweepg_c4_$TAG -f $CMD4 > $OUT/ncu_sweepg_$TAG.log 2>&1; echo "c4 capture exit $?"
ncu -i $OUT/sweepg_c4_$TAG.ncu-rep --page raw --csv > $OUT/sweepg_c4_${TAG}_raw.csv 2>/dev/null
LSMC_TRACE=$OUT/sweep_trace_$TAG.bin python bench.py --config c4 --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-panels > /dev/null 2>&1
python scripts/sweep_trace_report.py $OUT/sweep_trace_$TAG.bin > $OUT/sweep_trace_$TAG.txt; tail -5 $OUT/sweep_trace_$TAG.txt
find $OUT -name "*.ncu-rep" -size +14M -delete
