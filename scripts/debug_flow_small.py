"""The small flow-mode trainer case of tests/test_gpu_trainer.py with the counters printed after every step
(which step / how many waits time out), for A/B runs under different AINIC_GEMM_* settings."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("AINIC_TIMEOUT_MS", "1000")
os.environ.setdefault("AINIC_MLP_GRAPH", "0")
os.environ["AINIC_FLOW"] = "1"
from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer  # noqa: E402
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402

sizes = [256, 384, 264, 256, 128]
cfg = MLPConfig(sizes=sizes, mb=256, fuse=0, lr=0.5, loss_weight=1.0, bucket_mb=0.0)
g = torch.Generator().manual_seed(31)
xs = [K.cast_bf16((torch.rand(cfg.mb, sizes[0], generator=g) - 0.5).cuda()) for _ in range(2)]
ys = [torch.randint(0, sizes[-1], (cfg.mb,), generator=g, dtype=torch.int32).cuda() for _ in range(2)]
tr = MLPTrainer(cfg)
print({k: os.environ.get(k) for k in ("AINIC_GEMM_TMA_STORE", "AINIC_GEMM_PDL", "AINIC_MLP_GRAPH", "AINIC_FLOW_TRACE")})
for i in range(4):
    t = time.time()
    tr.step(xs[i % 2], ys[i % 2])
    tr.wait()
    torch.cuda.synchronize()
    c = tr.comm.counters()
    print(f"step {i}: {1e3 * (time.time() - t):.1f} ms  timeouts {c['timeouts']} launches {c['launches']}", flush=True)
    if os.environ.get("AINIC_FLOW_TRACE"):
        print("   trace", tr.comm.flow_trace(cfg.nlayers))
