#!/bin/bash
# Round-2 call r (1 GPU): 88-register GEMM (flow co-residency), timelines of the 1-CTA and 2-CTA kernels, tests, bench.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 120 python scripts/debug_flow_small.py > gpurun_out/r_flow_default.log 2>&1; tail -5 gpurun_out/r_flow_default.log
timeout 300 python scripts/gemm_timeline.py NT NT:256 NT:-256 NT:-208 NN:-256 TN:-256 NT:0:1792:8192:8192 NT:-256:1792:8192:8192 > gpurun_out/r_timeline.log 2>&1; echo "timeline rc=$?"; grep -v "^launch [0-46-7]" gpurun_out/r_timeline.log
timeout 300 python scripts/gemm_epi_bench.py --out gpurun_out/r_gemm_epi_tma.json > gpurun_out/r_gemm_epi_tma.log 2>&1; echo "rc=$?"; grep ours gpurun_out/r_gemm_epi_tma.log
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_trainer.py -m gpu -x -q > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r_pytest.log | cut -c1-300
for t in 1 0; do
  AINIC_GEMM_TMA_STORE=$t timeout 300 python bench.py --steps 20 --warmup 5 --no-baseline > gpurun_out/r_bench_tma${t}.json 2> gpurun_out/r_bench_tma${t}.err; echo "bench tma=$t rc=$?"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/r_bench_tma${t}.json').read().strip().splitlines()[-1]); print(d.get('ms_per_step'), 'e2e', d['e2e'].get('ms_per_step'), 'loss', d.get('loss'))"
done
