#!/bin/bash
# Round-2 call q (1 GPU): flow-mode small case under different GEMM settings; per-CTA GEMM timeline; GEMM numbers
# of the x16 double-buffered TMA epilogue; kernel tests.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 120 python scripts/debug_flow_small.py > gpurun_out/q_flow_default.log 2>&1; tail -6 gpurun_out/q_flow_default.log
AINIC_GEMM_TMA_STORE=0 timeout 120 python scripts/debug_flow_small.py > gpurun_out/q_flow_notma.log 2>&1; tail -6 gpurun_out/q_flow_notma.log
AINIC_GEMM_PDL=0 timeout 120 python scripts/debug_flow_small.py > gpurun_out/q_flow_nopdl.log 2>&1; tail -6 gpurun_out/q_flow_nopdl.log
AINIC_FLOW_TRACE=1 timeout 120 python scripts/debug_flow_small.py > gpurun_out/q_flow_trace.log 2>&1; tail -12 gpurun_out/q_flow_trace.log
timeout 200 python scripts/gemm_timeline.py > gpurun_out/q_timeline.log 2>&1; echo "timeline rc=$?"; cat gpurun_out/q_timeline.log
AINIC_GEMM_TMA_STORE=0 timeout 200 python scripts/gemm_timeline.py > gpurun_out/q_timeline_direct.log 2>&1; echo "timeline rc=$?"; cat gpurun_out/q_timeline_direct.log
timeout 300 python scripts/gemm_epi_bench.py --out gpurun_out/q_gemm_epi_tma.json > gpurun_out/q_gemm_epi_tma.log 2>&1; echo "rc=$?"; grep ours gpurun_out/q_gemm_epi_tma.log
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/q_pytest.log | cut -c1-300
for t in 1 0; do
  AINIC_GEMM_TMA_STORE=$t timeout 300 python bench.py --steps 20 --warmup 5 --no-baseline > gpurun_out/q_bench_tma${t}.json 2> gpurun_out/q_bench_tma${t}.err; echo "bench tma=$t rc=$?"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/q_bench_tma${t}.json').read().strip().splitlines()[-1]); print(d.get('ms_per_step'), 'e2e', d['e2e'].get('ms_per_step'), 'loss', d.get('loss'))"
done
