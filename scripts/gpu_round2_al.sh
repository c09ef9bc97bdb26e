#!/bin/bash
# Round-2 call al (2 GPUs): multi-rank paths after the heap-allocator change (collective close, retired staging).
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
export AINIC_TIMEOUT_MS=5000
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $RUN --master-port 29761 bench.py --gpus 2 --steps 20 --warmup 5 --no-baseline > gpurun_out/al_bench_n2.json 2> gpurun_out/al_bench_n2.err; echo "bench rc=$?"
timeout 400 $RUN --master-port 29762 scripts/mp_check.py > gpurun_out/al_mp_check_n2.log 2>&1; echo "mp_check rc=$?"; grep -i "summary" gpurun_out/al_mp_check_n2.log | cut -c1-300 | tail -2
python - <<PY
import json
d=json.loads(open('gpurun_out/al_bench_n2.json').read().strip().splitlines()[-1])
print('bench', d.get('ms_per_step'), 'e2e', d['e2e'].get('ms_per_step'), d.get('correctness'))
PY
tail -n 3 gpurun_out/al_bench_n2.err | cut -c1-300
