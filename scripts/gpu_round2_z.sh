#!/bin/bash
# Round-2 call z (1 GPU): final validation - whole GPU test tier, smoke, the bench as the driver runs it, ncu of the bf16 GEMMs.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/z_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/z_pytest.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/z_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/z_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/z_bench_n1.json 2> gpurun_out/z_bench_n1.err; echo "bench rc=$?"
timeout 120 python bench.py --impl reference > gpurun_out/z_bench_ref.json 2>&1; echo "ref rc=$?"; cat gpurun_out/z_bench_ref.json | cut -c1-300
timeout 120 python scripts/prof_gemm_kinds.py > gpurun_out/z_prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gemm_kernel" -c 3 -o gpurun_out/z_gemm -f python scripts/prof_gemm_kinds.py > gpurun_out/z_ncu.log 2>&1
echo "ncu gemm rc=$?"; tail -1 gpurun_out/z_ncu.log
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/z_bench_n1.json').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("ms_per_step","value","vs_baseline","dtype","gpu_launches")}, 'e2e', d['e2e'].get('ms_per_step'))
except Exception as e: print("ERR",e)
PY
