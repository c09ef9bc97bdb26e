#!/bin/bash
# Round-2 call m (1 GPU): the whole GPU test tier, smoke, the bench as the driver runs it (+ f32 row), GEMM kind
# throughput, ncu of the GEMM kinds and of the all-reduce kernels (2 ranks emulated on the one GPU).
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=5000
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/m_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/m_pytest.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/m_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/m_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/m_bench_n1.json 2> gpurun_out/m_bench_n1.err; echo "bench rc=$?"
timeout 300 python bench.py --steps 20 --warmup 5 --dtype f32 --no-baseline > gpurun_out/m_bench_n1_f32.json 2> gpurun_out/m_bench_n1_f32.err; echo "f32 bench rc=$?"
timeout 300 python scripts/gemm_kinds_bench.py > gpurun_out/m_gemm_kinds.log 2>&1; echo "kinds bench rc=$?"
timeout 120 python scripts/prof_gemm_kinds.py > gpurun_out/m_prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gemm_kernel|gemm_kind_kernel" -c 9 -o gpurun_out/m_gemm -f python scripts/prof_gemm_kinds.py > gpurun_out/m_ncu.log 2>&1
echo "ncu gemm rc=$?"; tail -1 gpurun_out/m_ncu.log
timeout 120 python scripts/prof_flow.py 10 > gpurun_out/m_flow_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"flow_emul|emul_ov" -c 2 -o gpurun_out/m_allreduce -f python scripts/prof_flow.py 10 > gpurun_out/m_ncu2.log 2>&1
echo "ncu allreduce rc=$?"; tail -1 gpurun_out/m_ncu2.log
for f in gpurun_out/m_bench_n1*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], {k:d.get(k) for k in ("ms_per_step","value","vs_baseline","dtype","gpu_launches")}, 'e2e', d.get('e2e') and d['e2e'].get('ms_per_step'))
except Exception as e: print("ERR",e)
PY
done
