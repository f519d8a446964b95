#!/bin/bash
# On an 8-GPU box: the strong-scaling bench lines at N = 8 (C3 1M paths in total, C4 4M, C5 2M) and N = 4 (C3).
OUT=gpurun_out; mkdir -p $OUT; TAG=$1
run() { n=$1; name=$2; shift 2
  LSMC_TAIL_TRACE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n "$@" > $OUT/bench_${name}_n${n}_$TAG.json 2> $OUT/bench_${name}_n${n}_$TAG.err
  python -c "
import json;d=json.load(open('$OUT/bench_${name}_n${n}_$TAG.json'));print('$name n$n', 'ms/step',round(d['ms_per_step'],3),{k:round(v,3) for k,v in d['phase_ms'].items()},'value %.4g'%d['value'],d['scaling'],'paths/gpu',d['config']['paths_per_gpu'],'overhead us',round(d['backward_overhead_us_per_period'],2),'e2e ms',round(d['e2e']['ms_per_step'],2),'npv',d['npv'])"
  grep "tail trace" $OUT/bench_${name}_n${n}_$TAG.err | tail -1
}
run 8 c3 --steps 5 --warmup 3 --no-panels
run 8 c4 --config c4 --steps 2 --warmup 1 --no-panels
run 8 c5 --config c5 --steps 2 --warmup 1 --no-panels
run 4 c3 --steps 5 --warmup 3 --no-panels
