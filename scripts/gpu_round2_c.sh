#!/bin/bash
# Round-2 third call (2 GPUs): process-per-GPU correctness, then flow / fused-epilogue / prefetch comparisons.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=3000
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 scripts/mp_check.py > gpurun_out/c_mp_check_n$N.log 2>&1; echo "mp_check rc=$?"; tail -6 gpurun_out/c_mp_check_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 scripts/scale_probe.py --out gpurun_out/c_probe_n$N.json --configs \
  'AINIC_FLOW=1' 'AINIC_FLOW=0' 'AINIC_FLOW=1,AINIC_FUSE_COMPRESS=1' 'AINIC_FLOW=0,AINIC_FUSE_COMPRESS=1' \
  'AINIC_FLOW=1,AINIC_PF_DEPTH=0' 'AINIC_FLOW=1,AINIC_COMM_CTAS=74' 'AINIC_FLOW=1,AINIC_MLP_GRAPH=0' 'AINIC_FLOW=1,CODEC=f32' \
  > gpurun_out/c_probe_n$N.log 2>&1; echo "probe rc=$?"
python - <<'PY'
import json,glob
for f in glob.glob('gpurun_out/c_probe_n*.json'):
    for r in json.load(open(f)):
        if 'error' in r: print(r); continue
        b=min(r['reps'],key=lambda x:x['ms_per_step'])
        print(f"{r['config']:45s} best {r['best_ms']:.4f} all {[round(x['ms_per_step'],4) for x in r['reps']]} calls {b['calls']} busy {b['busy_us']:.0f} b0 {b['b0_us']:.0f} b1 {b['b1_us']:.0f} wait {b['wait_us']:.0f} same {r['replicas_identical']} to {b['timeouts']} loss {r['loss']:.4f}")
PY
