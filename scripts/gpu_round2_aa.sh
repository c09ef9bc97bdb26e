#!/bin/bash
# Round-2 call aa (1 GPU): MX inference forward - test and bench.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 600 python -m pytest tests/test_gpu_trainer.py tests/test_gpu_gemm_kinds.py -m gpu -x -q -k "forward_mx or gemm_kinds" > gpurun_out/aa_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/aa_pytest.log | cut -c1-300
timeout 600 python scripts/mx_forward_bench.py > gpurun_out/aa_mx_forward.log 2>&1; echo "rc=$?"; cat gpurun_out/aa_mx_forward.log | tail -5
timeout 200 python scripts/gemm_kinds_bench.py 2>&1 | grep "quantize"
