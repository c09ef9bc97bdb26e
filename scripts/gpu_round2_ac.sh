#!/bin/bash
# Round-2 call ac (1 GPU): CTA-pair GEMMs inside the training step: deep and reference configs, AINIC_GEMM_2CTA=0/1.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
for t in 0 1; do
  AINIC_GEMM_2CTA=$t timeout 300 python bench/mlp_configs.py --config deep --steps 10 --warmup 3 --out gpurun_out/ac_deep_2cta$t.json > gpurun_out/ac_deep_2cta$t.log 2>&1; echo "deep 2cta=$t rc=$?"; tail -1 gpurun_out/ac_deep_2cta$t.log | cut -c1-400
  AINIC_GEMM_2CTA=$t timeout 300 python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/ac_bench_2cta$t.json 2> gpurun_out/ac_bench_2cta$t.err; echo "bench 2cta=$t rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/ac_bench_2cta$t.json').read().strip().splitlines()[-1]); print(d.get('ms_per_step'), d.get('loss'))"
done
