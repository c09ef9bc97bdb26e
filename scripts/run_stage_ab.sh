#!/bin/bash
# A/B on one box: compressed epilogue with / without shared-memory staging, inside the training step.
cd "$(dirname "$0")/.."
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"
for st in 0 1 0 1; do
  AINIC_GEMM_STAGE_FRAMES=$st AINIC_FUSE_COMPRESS=1 timeout 200 $TR bench.py --gpus $N --steps 300 --warmup 10 --no-e2e 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('n$N fuse1 stage $st', round(d['ms_per_step'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
