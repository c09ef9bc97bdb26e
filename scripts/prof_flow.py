"""ncu target for the all-reduce kernels with 2 ranks emulated on one GPU (same device code as the
process-per-GPU path; "peer" memory is local, so NVLink latency is absent but every other cost is there):
the flow kernel over 10 layer-sized chunks and the one-shot kernel over the same 10 layers in two calls.
Run plain first, then under ncu with -k regex:flow_emul / regex:emul_ov."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402

layers = int(sys.argv[1]) if len(sys.argv) > 1 else 10
n = 2048 * 2048
grads = torch.randn(2, layers * n, device="cuda") * 1e-3
ws = torch.randn(2, layers * n, device="cuda")
for rep in range(2):
    torch.cuda.synchronize()
    t0 = time.time()
    K.allreduce_flow_emulate(grads, ws, 0.1, K.CODEC_BFP, [n] * layers, K.CLEAN, want_grads=False, ctas_per_rank=73)
    torch.cuda.synchronize()
    t1 = time.time()
    K.allreduce_emulate(grads[:, : 5 * n].contiguous(), ws[:, : 5 * n].contiguous(), 0.1, K.CODEC_BFP, K.ALGO_ONESHOT, K.CLEAN,
                        want_grads=False, ctas_per_rank=74)
    torch.cuda.synchronize()
    t2 = time.time()
    print(f"flow emulate (incl. codec + alloc) {1e3 * (t1 - t0):.2f} ms; one-shot emulate of 5 layers {1e3 * (t2 - t1):.2f} ms")
print("ok")
