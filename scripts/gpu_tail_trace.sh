#!/bin/bash
# On the GPU box: phase stamps of one mid-horizon tail kernel (LSMC_TAIL_TRACE) at several sizes, bench lines at the 8-GPU
# per-GPU share, ncu captures of the forward / simulation kernels.  Usage: scripts/gpu_tail_trace.sh <tag>
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
for args in "--config c3" "--config c3 --paths-total 125000" "--config c4 --paths-total 62500" "--config c5 --paths-total 250000"; do
  echo "== $args"
  LSMC_TAIL_TRACE=1 python bench.py $args --steps 2 --warmup 1 --no-cpu-baseline --no-parity --no-panels 2> $OUT/tail_trace_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('ms/step',round(d['ms_per_step'],3),'phases',{k:round(v,3) for k,v in d['phase_ms'].items()},'sweep avg ms',round(d['roofline']['avg_launch_ms'],4),'overhead us/period',round(d['backward_overhead_us_per_period'],2))"
  grep "tail trace" $OUT/tail_trace_$TAG.err | tail -2
done
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-panels"
ncu --set full --clock-control none --import-source on -k regex:"forward2_kernel|simulate_pair_kernel" -c 2 -o $OUT/fwdsim2_$TAG -f $CMD > $OUT/ncu_fwdsim2_$TAG.log 2>&1
echo "fwd/sim capture exit $?"
ncu -i $OUT/fwdsim2_$TAG.ncu-rep --page raw --csv > $OUT/fwdsim2_${TAG}_raw.csv 2>/dev/null
ncu -i $OUT/fwdsim2_$TAG.ncu-rep --page source --csv --print-source sass > $OUT/fwdsim2_${TAG}_source.csv 2>/dev/null
find $OUT -name "*.ncu-rep" -size +14M -delete
