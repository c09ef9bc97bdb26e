#!/bin/bash
# Round-2 call i (N GPUs): per-layer time line of the flow kernel; one fused call per bucket for comparison.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=3000
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29582 scripts/scale_probe.py --repeats 2 --out gpurun_out/i_probe_n$N.json --configs \
  'AINIC_FLOW=1,AINIC_FLOW_TRACE=1' 'AINIC_FLOW=0' 'AINIC_FLOW=1,AINIC_FLOW_TRACE=1,AINIC_MLP_GRAPH=0' \
  > gpurun_out/i_probe_n$N.log 2>&1; echo "probe rc=$?"
python - <<'PY'
import json,glob
for f in glob.glob('gpurun_out/i_probe_n*.json'):
    for r in json.load(open(f)):
        if 'error' in r: print(r); continue
        b=min(r['reps'],key=lambda x:x['ms_per_step'])
        print(f"{r['config']:50s} best {r['best_ms']:.4f} calls {b['calls']} busy {b['busy_us']:.0f} b0 {b['b0_us']:.0f} b1 {b['b1_us']:.0f} same {r['replicas_identical']} to {b['timeouts']}")
        if r.get('flow_trace'):
            for k,v in r['flow_trace'].items(): print('   ',k.ljust(12),v)
PY
