#!/usr/bin/env python
"""The reference application's command line on B200 (sw/mlp_mpi_example_f32.cpp:270, :285-296):

    mlp_train.py iters MB fuse_type type bn bk bc C1 C2 ... CN
    torchrun --nproc-per-node 8 ... apps/mlp_train.py 20 14336 0 A 32 32 32 2048 2048 ... 2048

* ``MB`` is the GLOBAL minibatch, split evenly over the ranks (:301).
* ``type`` A = fused all-reduce + SGD through the engine (the live path, :686-790); B = the software
  baseline the NIC replaces: an NCCL all-reduce of every bucket followed by our SGD kernel
  (:590-680 used MPI_Iallreduce + libxsmm opt_exec); F = forward only (:543-588).
* ``bn bk bc`` are libxsmm blocking factors: accepted for compatibility, meaningless for tcgen05 tiles.
* env: AINIC_CODEC (auto|f32|bfp|mxfp8|bf16), AINIC_ALGO (auto|oneshot|twoshot|ring), AINIC_LR (0.1),
  DETAILED_PROFILE=1 (per-phase device times), CHECKPOINT=path (save at the end, resume if present),
  CHECK=1 (per-layer L1 norm of the weights and a bit comparison of the replicas after the run),
  AINIC_BUCKET_MB (fp32 MiB of gradients per fused call; 0 = one call per layer like the reference),
  AINIC_MLP_GRAPH=0 (direct launches instead of replaying the step as a CUDA graph),
  AINIC_FUSE_COMPRESS=0|1|2 (weight-gradient GEMM epilogue: fp32 / wire-format frames / frames stored
  into the reducing rank's memory).
Output: the reference's block (GFLOP / fp time / GFLOPS / PERFDUMP / profile lines, :794-816).
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer  # noqa: E402
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402
from fpga_ai_nic_b200.utils import checkpoint, profiling  # noqa: E402


def main(argv):
    if len(argv) > 1 and argv[1] in ("-h", "--help"):
        print(__doc__)
        return 0
    a = argv[1:]
    iters = int(a[0]) if len(a) > 0 else 10
    global_mb = int(a[1]) if len(a) > 1 else 32
    fuse = int(a[2]) if len(a) > 2 else 0
    typ = a[3] if len(a) > 3 else "A"
    sizes = [int(x) for x in a[7:]] if len(a) > 8 else [256, 256, 256]
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    assert global_mb % world == 0 and typ in ("A", "B", "F") and 0 <= fuse <= 3
    mb = global_mb // world
    cfg = MLPConfig(sizes=sizes, mb=mb, fuse=fuse, lr=float(os.environ.get("AINIC_LR", "0.1")),
                    codec=os.environ.get("AINIC_CODEC", "auto"), algo=os.environ.get("AINIC_ALGO", "auto"))
    if typ == "B":
        cfg.codec = "f32"
    tr = MLPTrainer(cfg)
    ck = os.environ.get("CHECKPOINT")
    start = 0
    if ck and os.path.exists(ck):
        start = checkpoint.load(tr, ck)["step"]
    # rank 0 generates, everyone receives its slice (MPI_Scatter of the activations, :453-455)
    gen = torch.Generator().manual_seed(1)
    x_all = torch.rand(global_mb, sizes[0], generator=gen) * 0.1 - 0.05
    x = K.cast_bf16(x_all[rank * mb:(rank + 1) * mb].contiguous().cuda())
    y = torch.zeros(mb, dtype=torch.int32, device="cuda")   # labels are never initialised in the reference (:412)
    detailed = os.environ.get("DETAILED_PROFILE") == "1"
    if rank == 0:
        print(f"MLP: iters {iters}, global MB {global_mb} ({world} ranks x {mb}), fuse {fuse}, type {typ}, layers {cfg.nlayers}")

    if typ == "B" and world > 1:
        # Software baseline: a per-rank engine with no peers and lr = 0 produces the gradients (the local
        # "update" is then a no-op); NCCL sums every bucket; our SGD kernel applies it.
        lr = cfg.lr
        tr.close()
        cfg_b = MLPConfig(sizes=sizes, mb=mb, fuse=fuse, lr=0.0, codec="f32", algo="oneshot")
        from fpga_ai_nic_b200.parallel.comm import Communicator
        tr = MLPTrainer(cfg_b, comm=Communicator(cfg_b.heap_bytes(), rank=0, world_size=1))

        def step():
            tr.step(x, y)
            tr.wait()
            for l in range(cfg.nlayers):
                g = tr.buffer("grad", l)
                dist.all_reduce(g)
                K.sgd_(tr.buffer("weight", l).view(-1), g.view(-1), lr, tr.buffer("shadow", l).view(-1))
    elif typ == "F":
        def step():
            tr.forward(x)
    else:
        typ = "A"          # B on one rank is A: there is nothing to all-reduce

        def step():
            tr.step(x, y)

    for _ in range(2):
        step()
    tr.wait()
    torch.cuda.synchronize()
    if detailed and typ == "A":
        tr.profile(True)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        step()
    tr.wait()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 1e3], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    seconds = float(t.item())
    phases = None
    if detailed and typ == "A":
        p = tr.profile(False)
        phases = {"fwd_fc": p["fwd_fc"], "fwd_loss": p["loss"], "bwd_loss": 0.0, "bwdupd_fc": p["bwdupd_fc"],
                  "bwdupd_first": p["bwdupd_fc"] / max(1, cfg.nlayers), "allreduce": p["update_tail"]}
    if typ != "F":
        tr.check_health()
    if rank == 0:
        print(profiling.perfdump(sizes, mb, world, iters, seconds, phases))
        if typ != "F":
            print(f"loss = {float(tr.loss_dev.item()) / mb:.6f}")
        c = tr.comm.counters()
        print(f"NIC latency (ms) = {c['busy_ns'] / 1e6:.3f}  stall = {c['wait_ns'] / 1e6:.3f}  timeouts = {c['timeouts']}")
    if os.environ.get("CHECK") == "1" and typ != "F":
        # The reference parses CHECK but only prints it (:227-228); its per-layer L1-norm check sits in
        # "#if 0" (:818-826).  Here: L1 norm of every layer's weights and a bit comparison of the replicas.
        same_all = True
        for l in range(cfg.nlayers):
            w = tr.buffer("weight", l)
            n1 = float(w.abs().sum(dtype=torch.float64).item())
            h = w.contiguous().view(torch.int32).to(torch.int64).sum().reshape(1)   # order-independent bit checksum
            same = True
            if world > 1:
                hs = [torch.empty_like(h) for _ in range(world)]
                dist.all_gather(hs, h)
                same = all(int(v.item()) == int(hs[0].item()) for v in hs)
            same_all = same_all and same
            if rank == 0:
                print(f"CHECK layer {l}: L1 norm of weights = {n1:.9g}, replicas {'identical' if same else 'DIFFER'}")
        if rank == 0:
            print(f"CHECK: weight replicas {'bit-identical' if same_all else 'differ'} across {world} ranks")
    if ck and rank == 0:
        checkpoint.save(tr, ck, step=start + iters)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
