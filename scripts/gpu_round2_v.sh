#!/bin/bash
# Round-2 call v (N GPUs): correctness across processes (mp_check incl. NVLS bf16), the sweep with the NVLS bf16 columns,
# the bench as the driver runs it, A/B of the TMA-store epilogues inside the multi-GPU step.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
export AINIC_TIMEOUT_MS=5000
N=${1:-2}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $RUN --master-port 29701 scripts/mp_check.py > gpurun_out/v_mp_check_n$N.log 2>&1; echo "mp_check rc=$?"; grep -i "nvls\|fail\|summary" gpurun_out/v_mp_check_n$N.log | cut -c1-300 | tail -8
timeout 600 $RUN --master-port 29702 bench/allreduce_sweep.py --min-kb 64 --max-mb 1024 --iters 6 --algos oneshot,twoshot --out gpurun_out/v_sweep_n$N.json > gpurun_out/v_sweep_n$N.log 2>&1; echo "sweep rc=$?"; tail -12 gpurun_out/v_sweep_n$N.log | cut -c1-400
timeout 600 $RUN --master-port 29703 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/v_bench_n$N.json 2> gpurun_out/v_bench_n$N.err; echo "bench rc=$?"
timeout 300 $RUN --master-port 29704 scripts/scale_probe.py --configs 'AINIC_GEMM_TMA_STORE=1' 'AINIC_GEMM_TMA_STORE=0' 'AINIC_GEMM_TMA_STORE=1' 'AINIC_GEMM_TMA_STORE=0' --repeats 2 --out gpurun_out/v_probe_n$N.json > gpurun_out/v_probe_n$N.log 2>&1; echo "probe rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/v_bench_n$N.json').read().strip().splitlines()[-1])
    print('bench', d.get('ms_per_step'), 'e2e', d['e2e'].get('ms_per_step'), 'vs', d.get('vs_baseline'), d.get('correctness'))
except Exception as e: print('ERR', e)
try:
    for r in json.load(open('gpurun_out/v_probe_n$N.json')): print(r.get('config'), r.get('best_ms'), r.get('error'), r.get('replicas_identical'))
except Exception as e: print('ERR', e)
PY
