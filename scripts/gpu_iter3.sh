#!/bin/bash
# On the GPU box: GPU test suite, then phase stamps of the tail kernel (LSMC_TAIL_TRACE) and bench lines for C3 / C3 at the
# 8-GPU share / C4.  Usage: scripts/gpu_iter3.sh <tag>
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest_gpu_$TAG.log
for args in "--config c3 --steps 5 --warmup 3" "--config c3 --paths-total 125000 --steps 5 --warmup 3" "--config c4 --paths-total 62500 --steps 2 --warmup 1" "--config c4 --steps 2 --warmup 1"; do
  echo "== $args"
  LSMC_TAIL_TRACE=1 python bench.py $args --no-cpu-baseline --no-panels 2> $OUT/iter3_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('ms/step',round(d['ms_per_step'],3),'phases',{k:round(v,3) for k,v in d['phase_ms'].items()},'sweep avg ms',round(d['roofline']['avg_launch_ms'],4),'overhead us/period',round(d['backward_overhead_us_per_period'],2),'npv',d['npv'],'parity',d.get('parity',{}).get('pass'))"
  grep "tail trace" $OUT/iter3_$TAG.err | tail -2
done
