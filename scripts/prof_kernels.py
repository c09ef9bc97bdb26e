"""Tiny driver for ncu captures: launches each hot kernel a few times on the reference's layer
shape (1792 x 2048 x 2048, 16 MiB buckets).  Run plain first, then under
`ncu --set full --clock-control none --import-source on -k regex:<name> -c 3 -o gpurun_out/<name>`.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402
from fpga_ai_nic_b200.parallel.comm import Communicator  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
mb, n, k = 1792, 2048, 2048
torch.manual_seed(0)
x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
w = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
dy = (torch.randn(mb, n, device="cuda") * 0.5).bfloat16()
reps = 3

if which in ("all", "gemm"):
    for _ in range(reps):
        K.gemm_fwd(x, w)
        K.gemm_bwd_data(dy, w)
        K.gemm_bwd_weight(dy, x)
if which in ("all", "comm"):
    count = n * k
    g = torch.randn(count, device="cuda") * 1e-3
    wf = torch.randn(count, device="cuda")
    sh = torch.empty(count, dtype=torch.bfloat16, device="cuda")
    for _ in range(reps):
        K.sgd_(wf, g, 0.1, sh)
        buf = K.bfp_compress(g, K.CLEAN)
        K.bfp_decompress(buf, count, K.CLEAN)
        K.mx_compress(g)
    comm = Communicator(256 << 20)
    gh = comm.empty(count)
    gh.copy_(g)
    for _ in range(reps):
        comm.allreduce_fused(gh, wf, 0.1, sh, codec=K.CODEC_F32, algo=K.ALGO_ONESHOT)
        comm.allreduce_fused(gh, wf, 0.1, sh, codec=K.CODEC_BFP, algo=K.ALGO_TWOSHOT)
    grads = torch.randn(4, 1 << 20, device="cuda") * 1e-3
    ws = torch.randn(4, 1 << 20, device="cuda")
    for _ in range(reps):
        K.allreduce_emulate(grads, ws, 0.1, K.CODEC_BFP, K.ALGO_TWOSHOT, K.CLEAN, ctas_per_rank=16)
    logits = torch.randn(mb, n, device="cuda").bfloat16()
    labels = torch.randint(0, n, (mb,), dtype=torch.int32, device="cuda")
    for _ in range(reps):
        K.softmax_xent(logits, labels)
torch.cuda.synchronize()
print("ok")
