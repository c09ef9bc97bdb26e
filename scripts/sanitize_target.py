"""Tiny workload for compute-sanitizer (ONE tool per gpurun call, B200_PROFILING.md):

    compute-sanitizer --tool memcheck  python scripts/sanitize_target.py
    compute-sanitizer --tool racecheck python scripts/sanitize_target.py comm
    compute-sanitizer --tool synccheck python scripts/sanitize_target.py gemm

Touches every kernel family once on small shapes: codecs, SGD, softmax, the three GEMM layouts
(1-CTA and 2-CTA), and a 4-rank emulated two-shot / ring all-reduce (the cross-rank flag protocol
runs inside one cooperative launch, so the tools see every rank's accesses).  The reference's
counterpart is its FIFO overflow/underflow checking and CDC constraints (hw/fifo.v:87-88,
hw/async_fifo.sdc) - hardware race protection rather than a software sanitizer.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
torch.manual_seed(0)
if which in ("all", "codec"):
    g = torch.randn(512 * 9 + 40, device="cuda") * 1e-3
    K.bfp_decompress(K.bfp_compress(g, K.HW_EXACT), g.numel(), K.HW_EXACT)
    K.mx_decompress(K.mx_compress(g), g.numel())
    w = torch.randn_like(g)
    K.sgd_(w, g, 0.1, torch.empty(g.numel(), dtype=torch.bfloat16, device="cuda"))
    logits = torch.randn(64, 256, device="cuda").bfloat16()
    K.softmax_xent(logits, torch.zeros(64, dtype=torch.int32, device="cuda"))
if which in ("all", "gemm"):
    x = (torch.randn(256, 256, device="cuda") * 0.5).bfloat16()
    wt = (torch.randn(512, 256, device="cuda") * 0.05).bfloat16()
    dy = (torch.randn(256, 512, device="cuda") * 0.5).bfloat16()
    for bn in (0, -256):
        K.gemm_fwd(x, wt, block_n=bn)
        K.gemm_bwd_data(dy, wt, block_n=bn)
        K.gemm_bwd_weight(dy, x, block_n=bn)
    K.gemm_bwd_weight(dy, x, compressed_out=torch.empty(K.bfp_bytes(512 * 256), dtype=torch.uint8, device="cuda"))
if which in ("all", "comm"):
    grads = torch.randn(4, 8192 * 4 + 512, device="cuda") * 1e-3
    ws = torch.randn_like(grads)
    for algo in (K.ALGO_TWOSHOT, K.ALGO_ONESHOT, K.ALGO_RING):
        K.allreduce_emulate(grads.clone(), ws.clone(), 0.1, K.CODEC_BFP, algo, K.CLEAN, ctas_per_rank=2)
torch.cuda.synchronize()
print("sanitize target ok")
