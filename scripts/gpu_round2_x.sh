#!/bin/bash
# Round-2 call x (N GPUs): mp_check with the corrected NVLS bf16 tolerance, bench as the driver runs it.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
export AINIC_TIMEOUT_MS=5000
N=${1:-4}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $RUN --master-port 29711 scripts/mp_check.py > gpurun_out/x_mp_check_n$N.log 2>&1; echo "mp_check rc=$?"; grep -i "nvls\|summary" gpurun_out/x_mp_check_n$N.log | cut -c1-300 | tail -4
timeout 600 $RUN --master-port 29713 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/x_bench_n$N.json 2> gpurun_out/x_bench_n$N.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/x_bench_n$N.json').read().strip().splitlines()[-1])
    print('bench', d.get('ms_per_step'), 'e2e', d['e2e'].get('ms_per_step'), 'vs', d.get('vs_baseline'), d.get('correctness'))
except Exception as e: print('ERR', e)
PY
