#!/bin/bash
# Round-2 call u (1 GPU): compressed-frame TMA epilogue: kernel + trainer tests, timeline, same-box A/B.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_trainer.py -m gpu -x -q > gpurun_out/u_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/u_pytest.log | cut -c1-300
timeout 200 python scripts/gemm_timeline.py TNC TN > gpurun_out/u_timeline_tnc.log 2>&1; echo "rc=$?"; grep -v "^launch [0-46-7]" gpurun_out/u_timeline_tnc.log
AINIC_GEMM_TMA_STORE=0 timeout 200 python scripts/gemm_timeline.py TNC > gpurun_out/u_timeline_tnc_direct.log 2>&1; echo "rc=$?"; grep -v "^launch [0-46-7]" gpurun_out/u_timeline_tnc_direct.log
