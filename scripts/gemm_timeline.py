"""Where the ~16 us of one reference-layer GEMM go: per-CTA time stamps written by the kernel itself
(ainic_gemm_trace: entry, after griddepcontrol.wait, first operand stage landed, last MMA issued, accumulator
complete, epilogue done, exit) for a chain of back-to-back launches, the way the training step issues them.
Prints, per launch, each stamp as (min / median / max over CTAs) in us after the first launch's first entry."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402

NAMES = ["entry", "dep_wait", "first_full", "mma_issued", "acc_done", "epi_done", "exit"]


def run(kind, bn=0, mb=1792, n=2048, k=2048, chain=8):
    torch.manual_seed(0)
    x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(n, k, device="cuda") * 0.02).bfloat16()
    dy = (torch.randn(mb, n, device="cuda") * 0.1).bfloat16()
    y = [torch.empty(mb, n, dtype=torch.bfloat16, device="cuda") for _ in range(2)]
    dw = torch.empty(n, k, dtype=torch.float32, device="cuda")
    cg = torch.empty(K.bfp_bytes(n * k), dtype=torch.uint8, device="cuda")

    def one(i):
        if kind == "NT":
            K.gemm_fwd(x if i == 0 else y[(i + 1) & 1], w, out=y[i & 1], block_n=bn)
        elif kind == "NN":
            K.gemm_bwd_data(dy if i == 0 else y[(i + 1) & 1], w, out=y[i & 1], block_n=bn)
        elif kind == "TNC":   # weight gradient emitted as BFP frames (what the trainer does from 2 ranks on)
            K.gemm_bwd_weight(dy, x, compressed_out=cg, block_n=bn)
        else:
            K.gemm_bwd_weight(dy, x, out=dw, block_n=bn)

    for _ in range(3):
        for i in range(chain):
            one(i)
    torch.cuda.synchronize()
    buf = torch.zeros(chain, 148, 8, dtype=torch.int64, device="cuda")
    K.gemm_trace(buf, chain)
    for i in range(chain):
        one(i)
    torch.cuda.synchronize()
    K.gemm_trace(None, 0)
    t = buf.cpu().double()
    live = t[:, :, 0] > 0
    t0 = t[0, :, 0][live[0]].min()
    print(f"--- {kind} block_n={bn} (negative = 2-CTA) {mb}x{n}x{k}: chain of {chain}; us after the first entry of launch 0 (min / median / max over CTAs)")
    for j in range(chain):
        m = live[j]
        if not m.any():
            continue
        row = [f"launch {j} ({int(m.sum())} CTAs)"]
        for s, name in enumerate(NAMES):
            v = (t[j, :, s][m] - t0) / 1e3
            v = v[v > -1e6]
            row.append(f"{name} {v.min():.1f}/{v.median():.1f}/{v.max():.1f}")
        print("  ".join(row))
    ex = [(t[j, :, 6][live[j]].max() - t0) / 1e3 for j in range(chain)]
    per = [(ex[j] - ex[j - 1]).item() for j in range(1, chain)]
    print("exit-to-exit per launch (us):", " ".join(f"{p:.1f}" for p in per))


if __name__ == "__main__":
    for spec in (sys.argv[1:] or ["NT", "NN", "TN"]):   # KIND[:block_n[:mb:n:k]]
        f = spec.split(":")
        dims = [int(v) for v in f[2:5]] if len(f) >= 5 else [1792, 2048, 2048]
        run(f[0], int(f[1]) if len(f) > 1 else 0, *dims)
