"""Time the weight-gradient GEMM with the fp32 epilogue vs the compressed (BFP / MXFP8) epilogue."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K
mb, n, k = 1792, 2048, 2048
dy = (torch.randn(mb, n, device="cuda") * 0.05).bfloat16()
x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
out = torch.empty(n, k, device="cuda")
cb = torch.empty(K.bfp_bytes(n * k), dtype=torch.uint8, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def t(fn, cold):
    for _ in range(5): fn()
    ts = []
    for _ in range(20):
        if cold: flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
    ts.sort(); return ts[len(ts) // 2] * 1e3
for name, fn in [("fp32", lambda: K.gemm_bwd_weight(dy, x, out=out)),
                 ("bfp_clean", lambda: K.gemm_bwd_weight(dy, x, compressed_out=cb, codec=K.CODEC_BFP, bfp_mode=K.CLEAN)),
                 ("bfp_hw", lambda: K.gemm_bwd_weight(dy, x, compressed_out=cb, codec=K.CODEC_BFP, bfp_mode=K.HW_EXACT)),
                 ("mxfp8", lambda: K.gemm_bwd_weight(dy, x, compressed_out=cb, codec=K.CODEC_MXFP8))]:
    print(f"{name:10s} cold {t(fn, True):6.1f} us  warm {t(fn, False):6.1f} us", flush=True)
