#!/bin/bash
# On the GPU box: one bench line per variant library built by scripts/build_variants.sh.  Usage: scripts/run_variants.sh name1 name2 ...
for n in "$@"; do
  LIB=$PWD/scratch_ab/liblsmc_$n.so; [ $n == base ] && LIB=$PWD/storage_b200/liblsmc_b200.so
  LSMC_B200_LIB=$LIB timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --paths-total ${PATHS:-1000000} --no-parity --no-panels --config ${CFG:-c3} 2>/dev/null |
    python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('variant $n: ms/step %.2f sweep %.4f ms  phases %s npv %.9f' % (d['ms_per_step'], d['roofline']['avg_launch_ms'], {k: round(v, 2) for k, v in d['phase_ms'].items()}, d['npv']))" || echo "variant $n failed"
done
