#!/bin/bash
# Round-2 second GPU call (1 GPU): flow-mode interlock after the shared-memory carve-out fix, rest of the GPU tier, DDP arms.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=2000
timeout 600 python -m pytest tests/test_gpu_trainer.py -m gpu -x -q -k flow > gpurun_out/b_flow.log 2>&1; echo "flow tests rc=$?"; tail -4 gpurun_out/b_flow.log
timeout 1200 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_allreduce_emul.py --deselect tests/test_gpu_kernels.py > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/b_pytest.log
timeout 200 env AINIC_FLOW=1 python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/b_bench_n1_flow.json 2> gpurun_out/b_bench_n1_flow.err; echo "flow bench rc=$?"
timeout 200 env AINIC_FLOW=1 AINIC_MLP_GRAPH=0 python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/b_bench_n1_flow_direct.json 2> gpurun_out/b_bench_n1_flow_direct.err; echo "flow direct bench rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/b_bench_n1.json 2> gpurun_out/b_bench_n1.err; echo "bench rc=$?"
for f in gpurun_out/b_bench_n1*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("ms_per_step","value","vs_baseline","gpu_launches")}, d.get("e2e") and d["e2e"].get("ms_per_step"), d.get("comm_counters"))
    print(json.dumps(d.get("baseline"))[:1800])
except Exception as e: print("ERR",e)
PY
done
