#!/bin/bash
# On the GPU box: C4 / C3 with programmatic dependent launch on neither, one or both of the two edges of the backward loop.
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
for cfg in "c4 --steps 2 --warmup 1" "c3 --steps 3 --warmup 2" "c3 --paths-total 125000 --steps 3 --warmup 2"; do
for edges in 3 1 2 0; do
  LSMC_PDL_EDGES=$edges python bench.py --config $cfg --no-cpu-baseline --no-panels --no-parity 2> $OUT/iter5_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$cfg edges $edges ms/step',round(d['ms_per_step'],3),'backward',round(d['phase_ms']['backward_ms'],2),'sweep avg ms',round(d['roofline']['avg_launch_ms'],4),'overhead us/period',round(d['backward_overhead_us_per_period'],2))"
done; done
