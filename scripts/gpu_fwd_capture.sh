#!/bin/bash
# On the GPU box: one ncu --set full capture of the streaming forward kernel (C3, 1M paths) with the source page.
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-panels"
ncu --set full --clock-control none --import-source on -k regex:"forward2_kernel" -c 1 -o $OUT/fwd2_$TAG -f $CMD > $OUT/ncu_fwd2_$TAG.log 2>&1
echo "fwd capture exit $?"
ncu -i $OUT/fwd2_$TAG.ncu-rep --page raw --csv > $OUT/fwd2_${TAG}_raw.csv 2>/dev/null
ncu -i $OUT/fwd2_$TAG.ncu-rep --page source --csv --print-source sass > $OUT/fwd2_${TAG}_source.csv 2>/dev/null
find $OUT -name "*.ncu-rep" -size +14M -delete
