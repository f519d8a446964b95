#!/bin/bash
# On an N-GPU box: backward-pass time of the C3 bench under different NCCL settings (the per-period all-reduce of the
# normal-equation block, 14 KB, sits on the critical path).  Usage: scripts/nccl_env_sweep.sh <n_gpus>
N=${1:-8}
run() {
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps 2 --warmup 2 --no-cpu-baseline 2>/dev/null |
    python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('%-60s ms/step %.2f backward %.2f forward %.2f' % ('$*', d['ms_per_step'], d['phase_ms']['backward_ms'], d['phase_ms']['forward_ms']))" || echo "$* failed"
}
run X=0
run NCCL_MAX_NCHANNELS=1
run NCCL_PROTO=LL
run NCCL_ALGO=Tree
run NCCL_ALGO=NVLS
run NCCL_ALGO=Ring NCCL_PROTO=LL NCCL_MAX_NCHANNELS=1
