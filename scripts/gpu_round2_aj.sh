#!/bin/bash
# Round-2 call aj (2 GPUs): the bench as the driver runs it with the 48 MiB default bucket; trainer checks across processes.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
export AINIC_TIMEOUT_MS=5000
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $RUN --master-port 29751 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/aj_bench_n2.json 2> gpurun_out/aj_bench_n2.err; echo "bench rc=$?"
timeout 600 $RUN --master-port 29752 scripts/mp_check.py > gpurun_out/aj_mp_check_n2.log 2>&1; echo "mp_check rc=$?"; grep -i "nvls\|summary" gpurun_out/aj_mp_check_n2.log | cut -c1-300 | tail -4
python - <<PY
import json
d=json.loads(open('gpurun_out/aj_bench_n2.json').read().strip().splitlines()[-1])
print('bench', d.get('ms_per_step'), 'e2e', d['e2e'].get('ms_per_step'), 'vs', d.get('vs_baseline'), d['config'].get('grad_bucket_mib'), d.get('correctness'))
PY
