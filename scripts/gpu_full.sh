#!/bin/bash
# On the GPU box: the whole GPU test suite, then bench lines C3 (1M and the 125k 8-GPU share), C4, C5 with tail phase stamps.
# Usage: scripts/gpu_full.sh <tag>
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest_gpu_$TAG.log
run() {
  name=$1; shift
  LSMC_TAIL_TRACE=1 python bench.py "$@" > $OUT/bench_${name}_$TAG.json 2> $OUT/bench_${name}_$TAG.err
  python -c "
import json;d=json.load(open('$OUT/bench_${name}_$TAG.json'));print('$name', 'ms/step',round(d['ms_per_step'],3),{k:round(v,3) for k,v in d['phase_ms'].items()},'sweep avg ms',round(d['roofline']['avg_launch_ms'],4),'frac',round(d['roofline']['frac'],4),'overhead us',round(d['backward_overhead_us_per_period'],2),'e2e ms',round(d['e2e']['ms_per_step'],2),'npv',d['npv'],'parity',d.get('parity',{}).get('pass'))"
  grep "tail trace" $OUT/bench_${name}_$TAG.err | tail -1
}
run c3 --steps 5 --warmup 3
run c3_125k --config c3 --paths-total 125000 --steps 5 --warmup 3 --no-cpu-baseline --no-panels
run c4 --config c4 --steps 2 --warmup 1 --no-cpu-baseline --no-panels
run c5 --config c5 --steps 2 --warmup 1 --no-cpu-baseline --no-panels
