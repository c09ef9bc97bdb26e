#!/bin/bash
# Round-2 call ai (N GPUs): SM budget of the all-reduce calls and bucket size, one process group.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
export AINIC_TIMEOUT_MS=5000
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741 scripts/scale_probe.py --repeats 2 \
  --configs AINIC_FLOW=0 AINIC_BUCKET_MB=32 AINIC_BUCKET_MB=48 AINIC_BUCKET_MB=64 AINIC_FLOW=0 AINIC_BUCKET_MB=48 \
  --out gpurun_out/ai_probe_n$N.json > gpurun_out/ai_probe_n$N.log 2>&1; echo "probe rc=$?"
python - <<PY
import json
for r in json.load(open('gpurun_out/ai_probe_n$N.json')): print(r.get('config'), r.get('best_ms'), [round(x['ms_per_step'],4) for x in r.get('reps',[])], r.get('error'), r.get('replicas_identical'))
PY
