#!/bin/bash
# Round-2 call g (N GPUs): ring-pipelined reduce stream - bit-exactness on one GPU, then step time and sweep at N.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=3000
N=${1:-2}
CUDA_VISIBLE_DEVICES=0 timeout 900 python -m pytest tests/test_gpu_allreduce_emul.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/g_emul.log 2>&1; echo "emul rc=$?"; tail -3 gpurun_out/g_emul.log | cut -c1-200
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29562 scripts/scale_probe.py --repeats 2 --out gpurun_out/g_probe_n$N.json --configs \
  'AINIC_FLOW=0' 'AINIC_FLOW=1' 'AINIC_FLOW=1,AINIC_COMM_CTAS=64' 'AINIC_FLOW=0,AINIC_BUCKET_LAYERS=4:3:2:1' 'AINIC_FLOW=0,AINIC_BUCKET_MB=0' \
  > gpurun_out/g_probe_n$N.log 2>&1; echo "probe rc=$?"
python - <<'PY'
import json,glob
for f in glob.glob('gpurun_out/g_probe_n*.json'):
    for r in json.load(open(f)):
        if 'error' in r: print(r); continue
        b=min(r['reps'],key=lambda x:x['ms_per_step'])
        print(f"{r['config']:50s} best {r['best_ms']:.4f} all {[round(x['ms_per_step'],4) for x in r['reps']]} calls {b['calls']} busy {b['busy_us']:.0f} b0 {b['b0_us']:.0f} b1 {b['b1_us']:.0f} same {r['replicas_identical']} to {b['timeouts']} loss {r['loss']:.4f}")
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29563 bench/allreduce_sweep.py --min-kb 1024 --max-mb 256 --codecs bfp,f32 --algos oneshot,twoshot --iters 6 --out gpurun_out/g_sweep_n$N.json > gpurun_out/g_sweep_n$N.log 2>&1; echo "sweep rc=$?"; tail -8 gpurun_out/g_sweep_n$N.log | cut -c1-220
