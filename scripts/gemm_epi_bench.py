"""bf16 GEMMs of the MLP (forward NT, backward-data NN, backward-weight TN) at the reference layer and
at two larger shapes, back-to-back launches inside one event pair, next to cuBLAS (torch.matmul).  Run once
with AINIC_GEMM_TMA_STORE=1 (default) and once with =0 to see what the TMA-store epilogue of a CTA's last
tile is worth.  Output: one JSON file named by --out."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402


def b2b(fn, n=40):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(n):
            fn()
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e) / n)
    return best


ap = argparse.ArgumentParser()
ap.add_argument("--out", default="gpurun_out/gemm_epi.json")
args = ap.parse_args()
rows = []
for (mb, n, k) in [(1792, 2048, 2048), (1792, 8192, 8192), (8192, 8192, 8192)]:
    torch.manual_seed(0)
    x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
    dy = (torch.randn(mb, n, device="cuda") * 0.1).bfloat16()
    y = torch.empty(mb, n, dtype=torch.bfloat16, device="cuda")
    dx = torch.empty(mb, k, dtype=torch.bfloat16, device="cuda")
    dw = torch.empty(n, k, dtype=torch.float32, device="cuda")
    impls = [
        ("NT fwd", "ours", lambda: K.gemm_fwd(x, w, out=y)), ("NT fwd", "cuBLAS", lambda: torch.matmul(x, w.t(), out=y)),
        ("NN dX", "ours", lambda: K.gemm_bwd_data(dy, w, out=dx)), ("NN dX", "cuBLAS", lambda: torch.matmul(dy, w, out=dx)),
        ("TN dW", "ours", lambda: K.gemm_bwd_weight(dy, x, out=dw)),
        ("TN dW", "cuBLAS bf16 out", lambda: torch.matmul(dy.t(), x)),
    ]
    for layout, impl, fn in impls:
        t = b2b(fn)
        r = {"layout": layout, "impl": impl, "mb": mb, "n": n, "k": k, "us": t * 1e3, "tflops": 2.0 * mb * n * k / t / 1e9}
        rows.append(r)
        print(f"{mb}x{n}x{k} {layout:7s} {impl:16s} {t*1e3:8.1f} us {r['tflops']:7.0f} TFLOPS", flush=True)
    # numerics of the path under test against fp32 PyTorch
    ref = x.float() @ w.float().t()
    err = ((K.gemm_fwd(x, w).float() - ref).norm() / ref.norm()).item()
    refw = dy.float().t() @ x.float()
    errw = ((K.gemm_bwd_weight(dy, x) - refw).norm() / refw.norm()).item()
    print(f"  rel err fwd {err:.2e}  dW {errw:.2e}", flush=True)
    rows.append({"mb": mb, "n": n, "k": k, "rel_err_fwd": err, "rel_err_dw": errw})
os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
json.dump({"tma_store": os.environ.get("AINIC_GEMM_TMA_STORE", "1"), "rows": rows}, open(args.out, "w"), indent=1)
