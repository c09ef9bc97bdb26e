#!/bin/bash
# Round-2 call f (N GPUs): flow v2 vs one fused call per bucket, worker-CTA counts, schedules; kinds GEMM tests on rank-0 GPU first.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=3000
N=${1:-2}
CUDA_VISIBLE_DEVICES=0 timeout 600 python -m pytest tests/test_gpu_gemm_kinds.py -m gpu -q > gpurun_out/f_kinds.log 2>&1; echo "kinds+trainer rc=$?"; tail -12 gpurun_out/f_kinds.log | cut -c1-200
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29552 scripts/scale_probe.py --repeats 2 --out gpurun_out/f_probe_n$N.json --configs \
  'AINIC_FLOW=1' 'AINIC_FLOW=0' 'AINIC_FLOW=1,AINIC_COMM_CTAS=64' \
  'AINIC_FLOW=1,CODEC=f32' \
  > gpurun_out/f_probe_n$N.log 2>&1; echo "probe rc=$?"; grep -v "^W1\|^$\|^\*\*\*\|OMP_NUM" gpurun_out/f_probe_n$N.log | grep -i "error\|Traceback" | head -5
python - <<'PY'
import json,glob
for f in glob.glob('gpurun_out/f_probe_n*.json'):
    for r in json.load(open(f)):
        if 'error' in r: print(r); continue
        b=min(r['reps'],key=lambda x:x['ms_per_step'])
        print(f"{r['config']:50s} best {r['best_ms']:.4f} all {[round(x['ms_per_step'],4) for x in r['reps']]} calls {b['calls']} busy {b['busy_us']:.0f} b0 {b['b0_us']:.0f} b1 {b['b1_us']:.0f} same {r['replicas_identical']} to {b['timeouts']} loss {r['loss']:.4f}")
PY
