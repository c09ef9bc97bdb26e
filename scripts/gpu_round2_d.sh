#!/bin/bash
# Round-2 fourth call (8 GPUs): bucket schedules / flow / prefetch at full world size, one process group.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=3000
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 scripts/scale_probe.py --repeats 2 --out gpurun_out/d_probe_n$N.json --configs \
  'AINIC_FLOW=0' 'AINIC_FLOW=0,AINIC_BUCKET_LAYERS=4:3:2:1' 'AINIC_FLOW=0,AINIC_BUCKET_LAYERS=3:3:2:1:1' 'AINIC_FLOW=0,AINIC_BUCKET_LAYERS=3:2:2:1:1:1' \
  'AINIC_FLOW=0,AINIC_BUCKET_MB=0' 'AINIC_FLOW=0,AINIC_PF_DEPTH=0' 'AINIC_FLOW=1,AINIC_FUSE_COMPRESS=1' 'AINIC_FLOW=0,AINIC_BUCKET_LAYERS=4:3:2:1,AINIC_FUSE_COMPRESS=0' \
  'AINIC_FLOW=0,AINIC_BUCKET_LAYERS=2:2:2:2:1:1' \
  > gpurun_out/d_probe_n$N.log 2>&1; echo "probe rc=$?"; tail -3 gpurun_out/d_probe_n$N.log | cut -c1-300
python - <<'PY'
import json,glob
for f in glob.glob('gpurun_out/d_probe_n*.json'):
    for r in json.load(open(f)):
        if 'error' in r: print(r); continue
        b=min(r['reps'],key=lambda x:x['ms_per_step'])
        print(f"{r['config']:62s} best {r['best_ms']:.4f} all {[round(x['ms_per_step'],4) for x in r['reps']]} calls {b['calls']} busy {b['busy_us']:.0f} b0 {b['b0_us']:.0f} b1 {b['b1_us']:.0f} wait {b['wait_us']:.0f} same {r['replicas_identical']} to {b['timeouts']}")
PY
