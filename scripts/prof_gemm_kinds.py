"""ncu target: the reference layer (1792 x 2048 x 2048) through the bf16, tf32 and MX block-scaled GEMMs,
forward (NT) and backward-weight (TN) layouts."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402

m, n, k = 1792, 2048, 2048
torch.manual_seed(0)
xf = torch.randn(m, k, device="cuda") * 0.5
wf = torch.randn(n, k, device="cuda") * 0.05
dyf = torch.randn(m, n, device="cuda") * 0.1
xb, wb, dyb = xf.bfloat16(), wf.bfloat16(), dyf.bfloat16()
xq, xs = K.mx_quantize_rows(xb)
wq, ws = K.mx_quantize_rows(wb)
dyq, dys = K.mx_quantize_cols(dyb)
xq2, xs2 = K.mx_quantize_cols(xb)
torch.cuda.synchronize()
K.gemm_fwd(xb, wb)
K.gemm_bwd_weight(dyb, xb)
K.gemm_bwd_data(dyb, wb)
K.gemm_tf32(K.GEMM_NT, xf, wf)
K.gemm_tf32(K.GEMM_TN, dyf, xf)
K.gemm_mxfp8(K.GEMM_NT, xq, xs, wq, ws, block_n=128)
K.gemm_mxfp8(K.GEMM_NT, xq, xs, wq, ws, block_n=256)
K.gemm_mxfp8(K.GEMM_TN, dyq, dys, xq2, xs2, out_dtype=torch.float32, block_n=128)
K.gemm_mxfp8(K.GEMM_TN, dyq, dys, xq2, xs2, out_dtype=torch.float32, block_n=256)
torch.cuda.synchronize()
print("ok")
