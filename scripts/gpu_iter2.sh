#!/bin/bash
# Development, on the GPU box: C3 parity tests, bench lines at 1M and 125k paths (the 8-GPU share), tail phase stamps.
# Usage: scripts/gpu_iter2.sh <tag> [pytest -k expression]
TAG=$1; KEXPR=${2:-"c3 or full_size or reference_fixture"}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "$KEXPR" > gpurun_out/pytest_iter_$TAG.log 2>&1; tail -3 gpurun_out/pytest_iter_$TAG.log
for args in "--config c3" "--config c3 --paths-total 125000"; do
  LSMC_TAIL_TRACE=1 python bench.py $args --steps 3 --warmup 2 --no-cpu-baseline --no-panels > gpurun_out/bench_iter_$TAG.json 2> gpurun_out/bench_iter_$TAG.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_iter_$TAG.json'));print('$args', 'ms/step',round(d['ms_per_step'],3),{k:round(v,3) for k,v in d['phase_ms'].items()},'sweep avg ms',round(d['roofline']['avg_launch_ms'],4),'overhead us',round(d['backward_overhead_us_per_period'],2),'e2e ms',round(d['e2e']['ms_per_step'],2),'npv',d['npv'],'parity',d.get('parity',{}).get('pass'))"
  grep "tail trace" gpurun_out/bench_iter_$TAG.err | tail -1
done
