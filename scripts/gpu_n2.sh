#!/bin/bash
# On a 2-GPU box: the whole GPU test suite (with the multi-GPU tests: cudaIpc exchange, NCCL fallback, groups in one process) and the
# strong-split C3 bench line at N = 2.
OUT=gpurun_out; mkdir -p $OUT; TAG=$1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_n2_$TAG.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_gpu_n2_$TAG.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 --no-panels > $OUT/bench_c3_n2_$TAG.json 2> $OUT/bench_c3_n2_$TAG.err
python -c "
import json;d=json.load(open('$OUT/bench_c3_n2_$TAG.json'));print('c3 n2', 'ms/step',round(d['ms_per_step'],3),{k:round(v,3) for k,v in d['phase_ms'].items()},'value %.4g'%d['value'],'overhead us',round(d['backward_overhead_us_per_period'],2),'e2e ms',round(d['e2e']['ms_per_step'],2),'npv',d['npv'],'parity',d.get('parity',{}).get('pass'))"
