"""ncu driver for the 2-CTA GEMM: a few launches of each layout with block_n = -256 (cta_group::2)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402

bn = int(sys.argv[1]) if len(sys.argv) > 1 else -256
mb, n, k = 1792, 2048, 2048
torch.manual_seed(0)
x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
w = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
dy = (torch.randn(mb, n, device="cuda") * 0.5).bfloat16()
for _ in range(3):
    K.gemm_fwd(x, w, block_n=bn)
    K.gemm_bwd_weight(dy, x, block_n=bn)
torch.cuda.synchronize()
print("ok")
