#!/bin/bash
# Round-2 call ad (1 GPU): where the end-to-end loop loses time on one GPU (diagnostic probes of bench.py).
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
for p in none noh2d noloss; do
  AINIC_BENCH_PROBE=$p timeout 300 python bench.py --steps 200 --warmup 10 --no-baseline > gpurun_out/ad_bench_$p.json 2> gpurun_out/ad_bench_$p.err; echo "probe=$p rc=$?"
  python -c "
import json; d=json.loads(open('gpurun_out/ad_bench_$p.json').read().strip().splitlines()[-1]); e=d['e2e']; print('$p', 'dev', d.get('ms_per_step'), 'e2e', e.get('ms_per_step'), 'bf16', e.get('bf16_host_inputs',{}).get('ms_per_step'))"
done
