#!/bin/bash
# Round-2 call ae (1 GPU): input upload as a kernel - test, end-to-end A/B against cudaMemcpyAsync + cast.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "upload" > gpurun_out/ae_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/ae_pytest.log | cut -c1-300
for i in 1 2; do for u in kernel dma; do
  AINIC_BENCH_UPLOAD=$u timeout 300 python bench.py --steps 100 --warmup 10 --no-baseline > gpurun_out/ae_bench_${u}_$i.json 2> gpurun_out/ae_bench_${u}_$i.err; echo "upload=$u rc=$?"
  python -c "
import json; d=json.loads(open('gpurun_out/ae_bench_${u}_$i.json').read().strip().splitlines()[-1]); e=d['e2e']; print('$u', 'dev', d.get('ms_per_step'), 'e2e', e.get('ms_per_step'), 'copy alone', e.get('h2d_copy_alone_ms'), 'loss', d.get('loss'))"
done; done
