"""Run every tcgen05 GEMM variant (layout x tile width) in its own subprocess, check it against a
PyTorch fp32 reference and time it on the reference's layer shape (1792 x 2048 x 2048,
sw/run.sh:16).  A faulting variant cannot poison the others; results go to
gpurun_out/gemm_matrix.json.

    python scripts/gpu_gemm_matrix.py            # parent: spawns the children
    python scripts/gpu_gemm_matrix.py child fwd 256
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# block_n > 0: 1-CTA kernel; < 0: 2-CTA (cta_group::2) kernel with tile width -block_n; 0: heuristic
VARIANTS = [("fwd", bn) for bn in (0, -256, -208, -128, 256, 208, 128)] + \
           [("bwd_data", bn) for bn in (0, -256, -128, 256, 192, 128)] + \
           [("bwd_weight", bn) for bn in (0, -256, -128, 256, 192, 128)]
SHAPES = [(128, 256, 64), (256, 512, 256), (384, 1024, 1088), (1792, 2048, 2048)]


def child(kind: str, bn: int) -> None:
    import torch
    from fpga_ai_nic_b200.ops import kernels as K
    res = []
    for (mb, n, k) in SHAPES:
        torch.manual_seed(0)
        x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
        w = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
        dy = (torch.randn(mb, n, device="cuda") * 0.5).bfloat16()
        if kind == "fwd":
            fn = lambda: K.gemm_fwd(x, w, block_n=bn)
            ref = x.float() @ w.float().t()
        elif kind == "bwd_data":
            fn = lambda: K.gemm_bwd_data(dy, w, block_n=bn)
            ref = dy.float() @ w.float()
        else:
            fn = lambda: K.gemm_bwd_weight(dy, x, block_n=bn)
            ref = dy.float().t() @ x.float()
        out = fn()
        torch.cuda.synchronize()
        err = (out.float() - ref).abs()
        tol = 0.02 * ref.abs().max().item() + 1e-3
        bad = err > tol
        entry = {"shape": [mb, n, k], "max_err": err.max().item(), "tol": tol, "bad_frac": bad.float().mean().item(),
                 "ok": not bool(bad.any())}
        if bad.any():
            rows = bad.any(1).nonzero().flatten()[:8].tolist()
            cols = bad.any(0).nonzero().flatten()[:16].tolist()
            entry["bad_rows"] = rows
            entry["bad_cols"] = cols
            entry["bad_rows_mod128"] = sorted(set(int(r) % 128 for r in bad.any(1).nonzero().flatten().tolist()))[:32]
            entry["bad_cols_mod64"] = sorted(set(int(c) % 64 for c in bad.any(0).nonzero().flatten().tolist()))[:64]
            entry["sample"] = [[out[r, c].item(), ref[r, c].item()] for r, c in zip(rows[:4], cols[:4])]
        if entry["ok"] and (mb, n, k) == SHAPES[-1]:
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
            ts = []
            for _ in range(10):
                flush.zero_()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                fn()
                e.record()
                torch.cuda.synchronize()
                ts.append(s.elapsed_time(e))
            ms = sorted(ts)[len(ts) // 2]
            entry["ms"] = ms
            entry["tflops"] = 2.0 * mb * n * k / ms / 1e9
            # cuBLAS reference point
            a, b = (x, w.t()) if kind == "fwd" else ((dy, w) if kind == "bwd_data" else (dy.t(), x))
            for _ in range(5):
                torch.matmul(a, b)
            ts = []
            for _ in range(10):
                flush.zero_()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                torch.matmul(a, b)
                e.record()
                torch.cuda.synchronize()
                ts.append(s.elapsed_time(e))
            entry["cublas_ms"] = sorted(ts)[len(ts) // 2]
        res.append(entry)
        print(json.dumps({"kind": kind, "bn": bn, **entry}), flush=True)
        if not entry["ok"]:
            break


def main() -> None:
    out = []
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    only = sys.argv[1:] if len(sys.argv) > 1 else None
    for kind, bn in VARIANTS:
        if only and kind not in only:
            continue
        t0 = time.time()
        try:
            p = subprocess.run([sys.executable, __file__, "child", kind, str(bn)], capture_output=True, text=True,
                               timeout=240)
            lines = [json.loads(l) for l in p.stdout.splitlines() if l.startswith("{")]
            out.append({"kind": kind, "bn": bn, "rc": p.returncode, "results": lines,
                        "stderr": p.stderr[-1500:] if p.returncode else ""})
        except subprocess.TimeoutExpired:
            out.append({"kind": kind, "bn": bn, "rc": "timeout", "results": []})
        last = out[-1]
        status = "OK" if last["rc"] == 0 and last["results"] and all(r["ok"] for r in last["results"]) else "FAIL"
        perf = next((f'{r["tflops"]:.0f} TFLOPS ({r["ms"]:.3f} ms, cuBLAS {r.get("cublas_ms", 0):.3f} ms)'
                     for r in last["results"] if "tflops" in r), "")
        print(f"{kind:11s} bn={bn:4d} {status} {perf} [{time.time() - t0:.0f}s]", flush=True)
        if status == "FAIL":
            print(json.dumps(last)[:3000], flush=True)
    with open(os.path.join(ROOT, "gpurun_out", "gemm_matrix.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "child":
        child(sys.argv[2], int(sys.argv[3]))
    else:
        main()
