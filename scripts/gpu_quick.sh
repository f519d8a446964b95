#!/bin/bash
# Development, on the GPU box: parity tests, one bench line at full size, pipeline trace of one sweep launch.
# Usage: scripts/gpu_quick.sh <tag> [notest]
TAG=$1
if [ "${2:-}" != "notest" ]; then python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; tail -3 gpurun_out/pytest_gpu_$TAG.log; fi
python bench.py --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/bench_$TAG.json 2>gpurun_out/bench_$TAG.err
python -c "
import json;d=json.load(open('gpurun_out/bench_$TAG.json'));print(d['ms_per_step'],d['phase_ms'],'sweep avg ms',d['roofline']['avg_launch_ms'],'npv',d['npv'])"
LSMC_TRACE=gpurun_out/trace_$TAG.bin python bench.py --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>&1
