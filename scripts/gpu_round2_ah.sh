#!/bin/bash
# Round-2 call ah (1 GPU): vectorised MX column quantiser - tests and timing.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
timeout 600 python -m pytest tests/test_gpu_gemm_kinds.py -m gpu -x -q > gpurun_out/ah_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/ah_pytest.log | cut -c1-300
python - <<'PY'
import torch, sys
sys.path.insert(0, '.')
from fpga_ai_nic_b200.ops import kernels as K
def t(fn, n=10):
    fn(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3
for (r, c) in ((1792, 2048), (1792, 8192), (8192, 8192)):
    x = torch.randn(r, c, device='cuda').bfloat16()
    print(r, c, 'rows %.1f us' % t(lambda: K.mx_quantize_rows(x)), 'cols %.1f us' % t(lambda: K.mx_quantize_cols(x)))
PY
