#!/bin/bash
# Round-2 call k (8 GPUs): bucket schedules at full world size in one process group, then the bench exactly as the driver runs it.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=3000
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29592 scripts/scale_probe.py --repeats 2 --out gpurun_out/k_probe_n$N.json --configs \
  'AINIC_FLOW=0' 'AINIC_BUCKET_LAYERS=4:3:2:1' 'AINIC_BUCKET_LAYERS=3:3:2:1:1' 'AINIC_BUCKET_LAYERS=5:3:1:1' 'AINIC_BUCKET_LAYERS=3:2:2:1:1:1' \
  'AINIC_BUCKET_MB=0' 'AINIC_FUSE_COMPRESS=0' 'AINIC_PF_DEPTH=0' 'AINIC_BUCKET_LAYERS=4:3:2:1,AINIC_COMM_CTAS=96' \
  > gpurun_out/k_probe_n$N.log 2>&1; echo "probe rc=$?"
python - <<'PY'
import json,glob
for f in glob.glob('gpurun_out/k_probe_n*.json'):
    for r in json.load(open(f)):
        if 'error' in r: print(r); continue
        b=min(r['reps'],key=lambda x:x['ms_per_step'])
        print(f"{r['config']:50s} best {r['best_ms']:.4f} all {[round(x['ms_per_step'],4) for x in r['reps']]} calls {b['calls']} busy {b['busy_us']:.0f} b0 {b['b0_us']:.0f} b1 {b['b1_us']:.0f} same {r['replicas_identical']} to {b['timeouts']}")
PY
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29593 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/k_bench_n$N.json 2> gpurun_out/k_bench_n$N.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/k_bench_n8.json').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("ms_per_step","value","vs_baseline","gpu_launches","step_ms")})
    print('e2e', d['e2e'] and {k:d['e2e'][k] for k in ('ms_per_step','value')}, 'correctness', d.get('correctness'))
    print(json.dumps(d.get('baseline'))[:1500])
except Exception as e: print("ERR",e)
PY
tail -5 gpurun_out/k_bench_n$N.err | cut -c1-300
