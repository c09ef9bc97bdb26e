#!/bin/bash
# Round-2 call ak (1 GPU): symmetric-heap allocator - trainer and all-reduce tests.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 900 python -m pytest tests/test_gpu_trainer.py tests/test_gpu_allreduce_emul.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/ak_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/ak_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1


