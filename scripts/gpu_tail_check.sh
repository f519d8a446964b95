#!/bin/bash
# On the GPU box: tail phase stamps at C4 for two path counts (is the K = 56 factorisation time a property of the kernel or of the clocks
# the sweep before it leaves behind?).  Usage: scripts/gpu_tail_check.sh
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
for n in 62500 500000; do
  (sleep 4; nvidia-smi --query-gpu=clocks.sm,power.draw --format=csv,noheader) &
  LSMC_TAIL_TRACE=1 python bench.py --config c4 --paths-total $n --steps 2 --warmup 1 --no-cpu-baseline --no-panels --no-parity 2>&1 >/dev/null | grep "tail trace" | head -1
  wait
done
