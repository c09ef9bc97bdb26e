#!/bin/bash
# Round-2 first GPU call (1 GPU): full GPU test tier, then the bench with its DDP arms, then flow / prefetch A-B runs.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=5000
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/a_bench_n1.json 2> gpurun_out/a_bench_n1.err; echo "bench rc=$?"
timeout 300 env AINIC_FLOW=1 python bench.py --steps 20 --warmup 5 --no-baseline > gpurun_out/a_bench_n1_flow.json 2> gpurun_out/a_bench_n1_flow.err; echo "flow rc=$?"
timeout 300 env AINIC_PF_DEPTH=0 python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/a_bench_n1_pf0.json 2> gpurun_out/a_bench_n1_pf0.err; echo "pf0 rc=$?"
timeout 300 env AINIC_PF_DEPTH=16 python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/a_bench_n1_pf16.json 2> gpurun_out/a_bench_n1_pf16.err; echo "pf16 rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/a_smoke.log 2>&1; echo "smoke rc=$?"
tail -3 gpurun_out/a_smoke.log
for f in gpurun_out/a_bench_n1*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("ms_per_step","value","vs_baseline","gpu_launches")}, d.get("e2e",{}) and d["e2e"].get("ms_per_step"), d.get("step_ms"))
    print(json.dumps(d.get("baseline"))[:1500])
except Exception as e: print("ERR",e)
PY
done
