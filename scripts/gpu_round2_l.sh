#!/bin/bash
# Round-2 call l (N GPUs): the bench exactly as the driver runs it (product + DDP arms + checks), the 1 KiB - 1 GiB sweep,
# and (N=4) the process-per-GPU correctness script.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=3000
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29602 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/l_bench_n$N.json 2> gpurun_out/l_bench_n$N.err; echo "bench rc=$?"
python - $N <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/l_bench_n{sys.argv[1]}.json').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("ms_per_step","value","vs_baseline","gpu_launches")})
    print('e2e', d['e2e'] and {k:d['e2e'][k] for k in ('ms_per_step','value')}, 'correctness', d.get('correctness'))
    b=d.get('baseline') or {}
    for m,v in (b.get('ddp') or {}).items(): print('  ddp',m,{k:v.get(k) for k in ('ms_per_step','e2e_ms_per_step','cuda_graph','error','cuda_graph_error')})
    print({k:b.get(k) for k in ('ratio_vs_naive','ratio_vs_hardened','e2e_ratio_vs_best')})
except Exception as e: print("ERR",e)
PY
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29603 bench/allreduce_sweep.py --min-kb 1 --max-mb 1024 --iters 6 --out gpurun_out/sweep_n$N.json > gpurun_out/l_sweep_n$N.log 2>&1; echo "sweep rc=$?"; tail -14 gpurun_out/l_sweep_n$N.log | cut -c1-330
if [ "$N" = "4" ]; then
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29604 scripts/mp_check.py > gpurun_out/mp_check_n$N.log 2>&1; echo "mp_check rc=$?"; tail -2 gpurun_out/mp_check_n$N.log
fi
