#!/bin/bash
# Round-2 call e (1 GPU): flow v2 (emulation + trainer interlock), tf32 / MX GEMMs, f32 trainer, N=1 bench rows.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=2000
timeout 900 python -m pytest tests/test_gpu_allreduce_emul.py -m gpu -x -q -k flow > gpurun_out/e_flow_emul.log 2>&1; echo "flow emul rc=$?"; tail -3 gpurun_out/e_flow_emul.log
timeout 600 python -m pytest tests/test_gpu_trainer.py -m gpu -x -q > gpurun_out/e_trainer.log 2>&1; echo "trainer rc=$?"; tail -3 gpurun_out/e_trainer.log
timeout 900 python -m pytest tests/test_gpu_gemm_kinds.py -m gpu -q > gpurun_out/e_kinds.log 2>&1; echo "kinds rc=$?"; tail -25 gpurun_out/e_kinds.log | cut -c1-200
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/e_noalloc.log 2>&1; echo "noalloc rc=$?"; tail -2 gpurun_out/e_noalloc.log
timeout 200 env AINIC_FLOW=1 python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/e_bench_n1_flow.json 2> gpurun_out/e_bench_n1_flow.err; echo "flow bench rc=$?"
timeout 200 python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e --dtype f32 > gpurun_out/e_bench_n1_f32.json 2> gpurun_out/e_bench_n1_f32.err; echo "f32 bench rc=$?"; tail -3 gpurun_out/e_bench_n1_f32.err
timeout 200 python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/e_bench_n1.json 2> gpurun_out/e_bench_n1.err; echo "bench rc=$?"
for f in gpurun_out/e_bench_n1*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("ms_per_step","value","dtype","gpu_launches","loss")}, d.get("comm_counters"))
except Exception as e: print("ERR",e)
PY
done
