#!/bin/bash
# Round-2 call am (2 GPUs): bench as the driver runs it after the last bench.py change.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29771 bench.py --gpus 2 --steps 20 --warmup 5 --no-baseline > gpurun_out/am_bench_n2.json 2> gpurun_out/am_bench_n2.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/am_bench_n2.json').read().strip().splitlines()[-1])
print('bench', d.get('ms_per_step'), 'e2e', d['e2e'].get('ms_per_step'), d.get('correctness'))
PY
