#!/bin/bash
# On the GPU box: average sweep launch time of each ablation build.
for n in "$@"; do
  LSMC_B200_LIB=$PWD/scratch_ab/liblsmc_ab$n.so timeout 120 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --paths-per-gpu ${PATHS:-200000} 2>/dev/null |
    python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ablate $n: sweep avg_launch_ms %.4f  backward_ms %.1f  frac %.3f' % (d['roofline']['avg_launch_ms'], d['phase_ms']['backward_ms'], d['roofline']['frac']))" || echo "ablate $n failed"
done
