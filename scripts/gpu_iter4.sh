#!/bin/bash
# On the GPU box: general-sweep variants.  GPU test suite with the in-tree library, the K > 16 parity cases with each variant
# library under scratch_ab/, C4 / C3 bench lines and the phase stamps of one C4 sweep per library.  Usage: scripts/gpu_iter4.sh <tag> [variants...]
TAG=$1; shift; OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest_gpu_$TAG.log
for v in base "$@"; do
  LIB=$PWD/scratch_ab/liblsmc_$v.so; [ $v == base ] && LIB=$PWD/storage_b200/liblsmc_b200.so
  echo "=== variant $v"
  if [ $v != base ]; then LSMC_B200_LIB=$LIB python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "k56 or k20 or c1 or terminal" 2>&1 | tail -2; fi
  for args in "--config c4 --steps 2 --warmup 1" "--config c3 --steps 3 --warmup 2"; do
    LSMC_B200_LIB=$LIB python bench.py $args --no-cpu-baseline --no-panels 2> $OUT/iter4_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$args', 'ms/step',round(d['ms_per_step'],3),'phases',{k:round(v,3) for k,v in d['phase_ms'].items()},'sweep avg ms',round(d['roofline']['avg_launch_ms'],4),'npv',d['npv'],'parity',d.get('parity',{}).get('pass'))"
  done
  LSMC_B200_LIB=$LIB LSMC_TRACE=$OUT/sweep_trace_${TAG}_$v.bin python bench.py --config c4 --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-panels > /dev/null 2>&1
  python scripts/sweep_trace_report.py $OUT/sweep_trace_${TAG}_$v.bin | tee $OUT/sweep_trace_${TAG}_$v.txt
done
