#!/bin/bash
# On the GPU box: parity tests, then tail phase stamps and bench lines at C3 (1M and 125k paths) and C4 (62 500 paths).
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
for cfg in "c3 --steps 3 --warmup 2" "c3 --paths-total 125000 --steps 3 --warmup 2" "c4 --paths-total 62500 --steps 2 --warmup 1"; do
  LSMC_TAIL_TRACE=1 python bench.py --config $cfg --no-cpu-baseline --no-panels --no-parity 2> $OUT/tailchk.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$cfg ms/step',round(d['ms_per_step'],3),'backward',round(d['phase_ms']['backward_ms'],2),'overhead us/period',round(d['backward_overhead_us_per_period'],2),'npv',d['npv'])"
  grep "tail trace" $OUT/tailchk.err | head -1
done
