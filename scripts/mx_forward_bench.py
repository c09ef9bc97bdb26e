"""Inference forward pass of the deep (8 x 8192 x 8192, 1792 rows) and reference MLPs: bf16 tensor cores
(MLPTrainer.forward) against the block-scaled fp8 kind (MLPTrainer.forward_mx: one quantiser launch per layer +
tcgen05.mma.kind::mxf8f6f4.block_scale GEMMs on pre-quantised weights).  Device-timed, 20 passes per event pair."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer  # noqa: E402
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402


def timed(fn, n=20):
    for _ in range(3):
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


rows = []
for name, (layers, width) in {"reference 10x2048": (10, 2048), "deep 8x8192": (8, 8192)}.items():
    cfg = MLPConfig.reference(layers, width, 1792)
    tr = MLPTrainer(cfg)
    g = torch.Generator().manual_seed(3)
    x = K.cast_bf16((torch.rand(cfg.mb, width, generator=g) * 0.1 - 0.05).cuda())
    ref = tr.forward(x).float().clone()
    mx = tr.forward_mx(x).float().clone()
    err = ((mx - ref).norm() / ref.norm()).item()
    t_bf16 = timed(lambda: tr.forward(x))
    t_mx = timed(lambda: tr.forward_mx(x))
    flops = 2.0 * cfg.mb * width * width * layers
    r = {"model": name, "bf16_ms": t_bf16, "mx_ms": t_mx, "speedup": t_bf16 / t_mx, "bf16_tflops": flops / t_bf16 / 1e9,
         "mx_tflops": flops / t_mx / 1e9, "logits_rel_l2_err_vs_bf16": err}
    rows.append(r)
    print(json.dumps(r), flush=True)
    tr.close()
    tr.comm.close()
    del tr
    torch.cuda.empty_cache()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/mx_forward.json", "w"), indent=1)
