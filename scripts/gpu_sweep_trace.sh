#!/bin/bash
# On the GPU box: phase stamps of one mid-horizon general sweep at C4 + a short bench line.  Usage: scripts/gpu_sweep_trace.sh <tag>
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
python bench.py --config c4 --steps 2 --warmup 1 --no-cpu-baseline --no-panels 2> $OUT/sweep_trace_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('c4 ms/step',round(d['ms_per_step'],3),'sweep avg ms',round(d['roofline']['avg_launch_ms'],4),'npv',d['npv'],'parity',d.get('parity',{}).get('pass'))"
LSMC_TRACE=$OUT/sweep_trace_$TAG.bin python bench.py --config c4 --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-panels > /dev/null 2>> $OUT/sweep_trace_$TAG.err
python scripts/sweep_trace_report.py $OUT/sweep_trace_$TAG.bin | tee $OUT/sweep_trace_$TAG.txt
