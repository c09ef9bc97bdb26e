#!/bin/bash
# Round-2 call w (1 GPU): 8 epilogue warps in the compressed weight-gradient kernel: tests, timeline A/B, emulated 2-rank step.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_trainer.py -m gpu -x -q > gpurun_out/w_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/w_pytest.log | cut -c1-300
timeout 200 python scripts/gemm_timeline.py TNC > gpurun_out/w_timeline_tnc8.log 2>&1; echo "rc=$?"; grep -v "^launch [0-46-7]" gpurun_out/w_timeline_tnc8.log
AINIC_GEMM_EPI_WARPS=4 timeout 200 python scripts/gemm_timeline.py TNC > gpurun_out/w_timeline_tnc4.log 2>&1; echo "rc=$?"; grep -v "^launch [0-46-7]" gpurun_out/w_timeline_tnc4.log
