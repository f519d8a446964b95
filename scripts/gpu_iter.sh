#!/bin/bash
# Development, on the GPU box: C3 parity tests + one bench line (no CPU baseline). Usage: scripts/gpu_iter.sh <tag> [config]
TAG=$1; CFG=${2:-c3}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "c3 or full_size" > gpurun_out/pytest_iter_$TAG.log 2>&1; tail -3 gpurun_out/pytest_iter_$TAG.log
python bench.py --config $CFG --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/bench_$TAG.json 2>gpurun_out/bench_$TAG.err
python -c "
import json;d=json.load(open('gpurun_out/bench_$TAG.json'));print(d['ms_per_step'],d['phase_ms'],'sweep avg ms',d['roofline']['avg_launch_ms'],'npv',d['npv'])"
