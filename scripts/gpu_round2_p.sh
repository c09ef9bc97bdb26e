#!/bin/bash
# Round-2 call p (1 GPU): TMA-store epilogue of the bf16 GEMMs: kernel tests, GEMM A/B (AINIC_GEMM_TMA_STORE=1/0),
# trainer tests, bench A/B/A/B.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=5000
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_trainer.py tests/test_gpu_gemm_kinds.py -m gpu -x -q > gpurun_out/p_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/p_pytest.log | cut -c1-300
timeout 300 python scripts/gemm_epi_bench.py --out gpurun_out/p_gemm_epi_tma.json > gpurun_out/p_gemm_epi_tma.log 2>&1; echo "rc=$?"; cat gpurun_out/p_gemm_epi_tma.log | grep -v cuBLAS
AINIC_GEMM_TMA_STORE=0 timeout 300 python scripts/gemm_epi_bench.py --out gpurun_out/p_gemm_epi_direct.json > gpurun_out/p_gemm_epi_direct.log 2>&1; echo "rc=$?"; cat gpurun_out/p_gemm_epi_direct.log
for i in 1 2; do
  for t in 1 0; do
    AINIC_GEMM_TMA_STORE=$t timeout 300 python bench.py --steps 20 --warmup 5 --no-baseline > gpurun_out/p_bench_tma${t}_$i.json 2> gpurun_out/p_bench_tma${t}_$i.err; echo "bench tma=$t rc=$?"
    python - gpurun_out/p_bench_tma${t}_$i.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d.get("ms_per_step"), 'e2e', d.get('e2e') and d['e2e'].get('ms_per_step'), 'loss', d.get('loss'))
except Exception as e: print("ERR",e)
PY
  done
done
