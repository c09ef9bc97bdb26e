#!/bin/bash
# Round-2 call j (1 GPU): ncu --set full of the flow kernel and the one-shot kernel (2 emulated ranks), plus GEMM kinds launch check.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=5000
python scripts/prof_flow.py 10 > gpurun_out/j_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"flow_emul|emul_ov" -c 2 -o gpurun_out/j_flow -f python scripts/prof_flow.py 10 > gpurun_out/j_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/j_plain.log; tail -3 gpurun_out/j_ncu.log
ls -la gpurun_out/j_flow.ncu-rep
