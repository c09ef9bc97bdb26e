#!/bin/bash
# Round-2 call s (1 GPU): timeline of the compressed weight-gradient epilogue (direct / staged), 2-CTA on the large shapes.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 200 python scripts/gemm_timeline.py TNC TN > gpurun_out/s_timeline_tnc.log 2>&1; echo "rc=$?"; grep -v "^launch [0-46-7]" gpurun_out/s_timeline_tnc.log
AINIC_GEMM_STAGE_FRAMES=1 timeout 200 python scripts/gemm_timeline.py TNC > gpurun_out/s_timeline_tnc_staged.log 2>&1; echo "rc=$?"; grep -v "^launch [0-46-7]" gpurun_out/s_timeline_tnc_staged.log
AINIC_GEMM_2CTA=1 timeout 300 python scripts/gemm_epi_bench.py --out gpurun_out/s_gemm_epi_2cta.json > gpurun_out/s_gemm_epi_2cta.log 2>&1; echo "rc=$?"; cat gpurun_out/s_gemm_epi_2cta.log
timeout 300 python scripts/gemm_epi_bench.py --out gpurun_out/s_gemm_epi_1cta.json > gpurun_out/s_gemm_epi_1cta.log 2>&1; echo "rc=$?"; grep ours gpurun_out/s_gemm_epi_1cta.log
