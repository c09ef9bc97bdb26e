#!/bin/bash
# Round-2 final check (1 GPU): whole GPU test tier, smoke, bench as the driver runs it.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/final_pytest.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/final_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/final_bench_n1.json 2> gpurun_out/final_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/final_bench_n1.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("ms_per_step","value","vs_baseline","dtype","gpu_launches")}, 'e2e', d['e2e'].get('ms_per_step'))
PY
