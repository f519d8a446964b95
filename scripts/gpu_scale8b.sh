#!/bin/bash
# On an 8-GPU box: strong-split bench lines of the closing build — C3 (1M paths in total) at N = 8 and 4, C4 (4M paths) at N = 8.
OUT=gpurun_out; mkdir -p $OUT; TAG=$1
run() { n=$1; name=$2; shift 2
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n "$@" > $OUT/bench_${name}_n${n}_$TAG.json 2> $OUT/bench_${name}_n${n}_$TAG.err
  python -c "
import json;d=json.load(open('$OUT/bench_${name}_n${n}_$TAG.json'));print('$name n$n', 'ms/step',round(d['ms_per_step'],3),{k:round(v,3) for k,v in d['phase_ms'].items()},'value %.4g'%d['value'],d['scaling'],'paths/gpu',d['config']['paths_per_gpu'],'overhead us',round(d['backward_overhead_us_per_period'],2),'e2e ms',round(d['e2e']['ms_per_step'],2),'npv',d['npv'])"
}
run 8 c3 --steps 5 --warmup 3 --no-panels --no-parity
run 4 c3 --steps 5 --warmup 3 --no-panels --no-parity
run 8 c4 --config c4 --steps 2 --warmup 1 --no-panels --no-parity
