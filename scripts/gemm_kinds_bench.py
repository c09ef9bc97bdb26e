"""Throughput of the three tcgen05 operand kinds on the MLP's shapes: bf16 (kind::f16), fp32-as-tf32
(kind::tf32) and MXFP8 block-scaled (kind::mxf8f6f4.block_scale), forward (NT) and backward-weight (TN)
layouts, against cuBLAS bf16 / tf32 (torch.matmul).  Back-to-back launches inside one event pair (what a
chain of layers sees).  Output: gpurun_out/gemm_kinds.json + a table."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402


def b2b(fn, n=20):
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n


rows = []
torch.backends.cuda.matmul.allow_tf32 = True
for (m, n, k) in [(1792, 2048, 2048), (1792, 8192, 8192), (8192, 8192, 8192)]:
    torch.manual_seed(0)
    xf = torch.randn(m, k, device="cuda") * 0.5
    wf = torch.randn(n, k, device="cuda") * 0.05
    xb, wb = xf.bfloat16(), wf.bfloat16()
    ob = torch.empty(m, n, dtype=torch.bfloat16, device="cuda")
    of = torch.empty(m, n, dtype=torch.float32, device="cuda")
    xq, xs = K.mx_quantize_rows(xb)
    wq, wsf = K.mx_quantize_rows(wb)
    impls = {
        "bf16 ours": lambda: K.gemm_fwd(xb, wb, out=ob),
        "bf16 cuBLAS": lambda: torch.matmul(xb, wb.t(), out=ob),
        "tf32 ours": lambda: K.gemm_tf32(K.GEMM_NT, xf, wf, out=of),
        "tf32 cuBLAS": lambda: torch.matmul(xf, wf.t(), out=of),
        "mxfp8 ours bn128": lambda: K.gemm_mxfp8(K.GEMM_NT, xq, xs, wq, wsf, out=ob, block_n=128),
        "mxfp8 ours bn256": lambda: K.gemm_mxfp8(K.GEMM_NT, xq, xs, wq, wsf, out=ob, block_n=256),
        "mx quantize rows (x)": lambda: K.mx_quantize_rows(xb),
    }
    for name, fn in impls.items():
        t = b2b(fn)
        r = {"impl": name, "layout": "NT", "m": m, "n": n, "k": k, "ms": t, "tflops": 2.0 * m * n * k / t / 1e9}
        rows.append(r)
        print(f"NT {m}x{n}x{k} {name:22s} {t*1e3:8.1f} us  {r['tflops']:7.0f} TFLOPS", flush=True)
    # backward weight: dW[n,k] = dY[mb,n]^T X[mb,k]   (K = mb)
    dyf = torch.randn(m, n, device="cuda") * 0.1
    dyb = dyf.bfloat16()
    dw = torch.empty(n, k, dtype=torch.float32, device="cuda")
    dyq, dys = K.mx_quantize_cols(dyb)
    xq2, xs2 = K.mx_quantize_cols(xb)
    impls = {"bf16 ours": lambda: K.gemm_bwd_weight(dyb, xb, out=dw),
             "bf16 cuBLAS": lambda: torch.matmul(dyb.t(), xb),
             "tf32 ours": lambda: K.gemm_tf32(K.GEMM_TN, dyf, xf, out=dw),
             "tf32 cuBLAS": lambda: torch.matmul(dyf.t(), xf, out=dw),
             "mxfp8 ours bn128": lambda: K.gemm_mxfp8(K.GEMM_TN, dyq, dys, xq2, xs2, out=dw, block_n=128),
             "mxfp8 ours bn256": lambda: K.gemm_mxfp8(K.GEMM_TN, dyq, dys, xq2, xs2, out=dw, block_n=256)}
    for name, fn in impls.items():
        t = b2b(fn)
        r = {"impl": name, "layout": "TN", "m": n, "n": k, "k": m, "ms": t, "tflops": 2.0 * m * n * k / t / 1e9}
        rows.append(r)
        print(f"TN {n}x{k}x{m} {name:22s} {t*1e3:8.1f} us  {r['tflops']:7.0f} TFLOPS", flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/gemm_kinds.json", "w"), indent=1)
