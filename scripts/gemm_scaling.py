"""How the tcgen05 GEMM time splits into fixed cost and per-K cost, and its asymptotic throughput.
Times the forward layout for K in {256..8192} at the reference's M x N (1792 x 2048) and one large
square problem, for the 1-CTA and 2-CTA kernels and cuBLAS (torch.matmul), warm L2 (no flush) and
cold L2 (256 MiB memset between launches).  Output: gpurun_out/gemm_scaling.json + a table."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, cold, iters=12):
    for _ in range(4):
        fn()
    ts = []
    for _ in range(iters):
        if cold:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def back_to_back(fn, n=20):
    """n launches in a row inside one event pair: what a chain of layers sees (PDL overlap included)."""
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
shapes = [(1792, 2048, k) for k in (256, 512, 1024, 2048, 4096, 8192)] + [(8192, 8192, 8192), (4096, 4096, 4096)]
for (m, n, k) in shapes:
    torch.manual_seed(0)
    x = (torch.randn(m, k, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
    out = torch.empty(m, n, dtype=torch.bfloat16, device="cuda")
    impls = {"1cta_256": lambda: K.gemm_fwd(x, w, out=out, block_n=256),
             "1cta_208": lambda: K.gemm_fwd(x, w, out=out, block_n=208),
             "2cta_256": lambda: K.gemm_fwd(x, w, out=out, block_n=-256),
             "2cta_208": lambda: K.gemm_fwd(x, w, out=out, block_n=-208),
             "cublas": lambda: torch.matmul(x, w.t(), out=out)}
    for name, fn in impls.items():
        r = {"impl": name, "m": m, "n": n, "k": k, "cold_ms": timeit(fn, True), "warm_ms": timeit(fn, False),
             "b2b_ms": back_to_back(fn)}
        r["tflops_b2b"] = 2.0 * m * n * k / r["b2b_ms"] / 1e9
        rows.append(r)
        print(f"{name:9s} {m}x{n}x{k}: cold {r['cold_ms']*1e3:7.1f} us  warm {r['warm_ms']*1e3:7.1f} us  "
              f"back-to-back {r['b2b_ms']*1e3:7.1f} us  ({r['tflops_b2b']:.0f} TFLOPS)", flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/gemm_scaling.json", "w"), indent=1)
