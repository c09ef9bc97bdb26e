"""The native MLP trainer on one GPU: every fuse mode against the CPU oracle of
sw/mlp_mpi_example_f32.cpp's math, checkpoint / resume, per-phase profile,
health check and the reference-compatible command line."""
import os
import re
import subprocess
import sys

import pytest
import torch

from fpga_ai_nic_b200.ops import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _oracle_steps(cfg, w_init, b, x, y, steps):
    """fp32 oracle with the same bf16 roundings the GPU path applies to activations / weights."""
    W = [w.clone() for w in w_init]
    bf = lambda t: t.bfloat16().float()
    losses = []
    for _ in range(steps):
        acts, masks = [bf(x)], []
        for l in range(cfg.nlayers):
            yl, m = O.fc_fwd(acts[-1], bf(W[l]), b[l], cfg.fuse)
            acts.append(bf(yl))
            masks.append(m)
        loss, _, d = O.softmax_loss(acts[-1], y, cfg.loss_weight)
        losses.append(loss)
        d = bf(d)
        for l in reversed(range(cfg.nlayers)):
            dx, dw, _ = O.fc_bwd(acts[l], bf(W[l]), d, masks[l], cfg.fuse, need_dx=l > 0)
            W[l] = O.sgd(W[l], dw, cfg.lr)
            if l > 0:
                d = bf(dx)
    return W, losses


@pytest.mark.parametrize("fuse", [0, 1, 2, 3])
def test_trainer_matches_oracle(fuse):
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    cfg = MLPConfig(sizes=[256, 384, 256, 128], mb=256, fuse=fuse, lr=0.5, loss_weight=1.0)
    tr = MLPTrainer(cfg)
    g = torch.Generator().manual_seed(11)
    x = torch.rand(cfg.mb, cfg.sizes[0], generator=g) * 2 - 1
    y = torch.randint(0, cfg.sizes[-1], (cfg.mb,), generator=g, dtype=torch.int32)
    w0 = [tr.buffer("weight", l).cpu().clone() for l in range(cfg.nlayers)]
    b = [tr.buffer("bias", l).cpu().clone() for l in range(cfg.nlayers)]
    xb, yd = K.cast_bf16(x.cuda()), y.cuda()
    losses = []
    for _ in range(3):
        tr.step(xb, yd)
        tr.wait()
        torch.cuda.synchronize()
        losses.append(float(tr.loss_dev.item()) / cfg.mb)
    W, ref_losses = _oracle_steps(cfg, w0, b, x, y, 3)
    for a, r in zip(losses, ref_losses):
        assert abs(a - r) <= 2e-2 * abs(r) + 1e-3, (losses, ref_losses)
    for l in range(cfg.nlayers):
        upd = (W[l] - w0[l]).norm().clamp_min(1e-12)
        assert ((tr.buffer("weight", l).cpu() - W[l]).norm() / upd).item() < 0.03
        # biases are never updated, as in the reference
        assert torch.equal(tr.buffer("bias", l).cpu(), b[l])
        # the bf16 shadow the GEMMs read is the rounded master copy
        assert torch.equal(tr.buffer("shadow", l).cpu(), tr.buffer("weight", l).cpu().bfloat16())
    assert tr.check_health()["timeouts"] == 0
    tr.close()


def test_checkpoint_resume_is_bit_identical(tmp_path):
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    from fpga_ai_nic_b200.utils import checkpoint
    cfg = MLPConfig(sizes=[256, 256, 128], mb=128, lr=0.5, loss_weight=1.0)
    g = torch.Generator().manual_seed(3)
    x = K.cast_bf16((torch.rand(cfg.mb, 256, generator=g) - 0.5).cuda())
    y = torch.randint(0, 128, (cfg.mb,), generator=g, dtype=torch.int32).cuda()

    def run(tr, n):
        for _ in range(n):
            tr.step(x, y)
        tr.wait()
        torch.cuda.synchronize()

    a = MLPTrainer(cfg)
    run(a, 2)
    path = str(tmp_path / "ck.pt")
    checkpoint.save(a, path, step=2)
    run(a, 2)
    want = [a.buffer("weight", l).cpu().clone() for l in range(cfg.nlayers)]
    a.close()
    b = MLPTrainer(cfg)
    assert checkpoint.load(b, path)["step"] == 2
    run(b, 2)
    for l in range(cfg.nlayers):
        assert torch.equal(b.buffer("weight", l).cpu(), want[l])
    b.close()


def test_detailed_profile_phases():
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    cfg = MLPConfig(sizes=[512, 512, 512], mb=256)
    tr = MLPTrainer(cfg)
    x = K.cast_bf16(torch.rand(cfg.mb, 512).cuda())
    y = torch.zeros(cfg.mb, dtype=torch.int32, device="cuda")
    tr.step(x, y)
    tr.profile(True)
    for _ in range(4):
        tr.step(x, y)
    p = tr.profile(False)
    assert p["steps"] == 4 and p["fwd_fc"] > 0 and p["bwdupd_fc"] > 0 and p["loss"] > 0
    tr.close()


def test_reference_command_line_on_gpu():
    env = dict(os.environ, DETAILED_PROFILE="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "apps", "mlp_train.py"), "5", "256", "3", "A", "32", "32", "32",
                        "256", "512", "256"], capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stderr[-2000:]
    assert re.search(r"^PERFDUMP,BP,[^,]+,1,256,256,512,[0-9.]+,[0-9.]+$", p.stdout, re.M), p.stdout
    assert "Bwdupd compute FIRST time overlaped" in p.stdout and "loss = " in p.stdout
    p = subprocess.run([sys.executable, os.path.join(ROOT, "apps", "mlp_train.py"), "3", "128", "0", "F", "32", "32", "32",
                        "256", "256"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and "PERFDUMP" in p.stdout


@pytest.mark.parametrize("bucket_mb", [0.2, 1000.0])
def test_bucketed_allreduce_same_bits_as_per_layer(bucket_mb):
    """Several layers per fused call (layer slots are padded to 1024 values inside one arena; 264 x 264
    is not a multiple of 1024, so the bucket runs across zero padding) == one call per layer (:754)."""
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    sizes = [264, 264, 136, 264, 128]
    g = torch.Generator().manual_seed(5)
    x = K.cast_bf16((torch.rand(128, sizes[0], generator=g) - 0.5).cuda())
    y = torch.randint(0, sizes[-1], (128,), generator=g, dtype=torch.int32).cuda()
    out = []
    for mb in (0.0, bucket_mb):
        tr = MLPTrainer(MLPConfig(sizes=sizes, mb=128, fuse=2, lr=0.5, loss_weight=1.0, bucket_mb=mb))
        assert abs(tr.bucket_mb - mb) < 1e-6
        n0 = tr.comm.counters()["launches"]
        for _ in range(3):
            tr.step(x, y)
        tr.wait()
        torch.cuda.synchronize()
        calls = tr.comm.counters()["launches"] - n0
        out.append(([tr.buffer("weight", l).cpu().clone() for l in range(4)], calls))
        tr.close()
    assert out[0][1] == 12 and out[1][1] == (3 if bucket_mb > 100 else 6)
    for a, b in zip(out[0][0], out[1][0]):
        assert torch.equal(a, b)


def test_graph_replay_same_bits_as_direct_launches(monkeypatch):
    """On one rank the step is captured into a CUDA graph per (input, labels, loss) tuple and replayed;
    results, launch counts and mixing with direct steps (profile mode) must not change anything."""
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    cfg = MLPConfig(sizes=[256, 384, 256, 128], mb=256, fuse=3, lr=0.5, loss_weight=1.0)
    g = torch.Generator().manual_seed(21)
    xs = [K.cast_bf16((torch.rand(cfg.mb, 256, generator=g) - 0.5).cuda()) for _ in range(2)]
    ys = [torch.randint(0, 128, (cfg.mb,), generator=g, dtype=torch.int32).cuda() for _ in range(2)]
    out = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("AINIC_MLP_GRAPH", mode)
        tr = MLPTrainer(cfg)
        l0 = tr.launches
        losses = []
        for i in range(6):
            if mode == "1" and i == 3:
                tr.profile(True)              # profile mode runs direct launches in between replays
            tr.step(xs[i % 2], ys[i % 2])
            if mode == "1" and i == 3:
                assert tr.profile(False)["steps"] == 1
            losses.append(tr.loss_dev.clone())
        tr.wait()
        torch.cuda.synchronize()
        out[mode] = ([tr.buffer("weight", l).cpu().clone() for l in range(cfg.nlayers)], [float(v) for v in losses],
                     tr.launches - l0)
        assert tr.check_health()["timeouts"] == 0
        tr.close()
    assert out["0"][2] == out["1"][2] > 0
    # the loss is an atomicAdd reduction over rows (order varies run to run); the weights do not depend on it
    assert out["0"][1] == pytest.approx(out["1"][1], rel=1e-5)
    for a, b in zip(out["0"][0], out["1"][0]):
        assert torch.equal(a, b)


@pytest.mark.parametrize("graph", ["0", "1"])
@pytest.mark.parametrize("fuse", [0, 3])
def test_flow_mode_same_bits_as_per_layer_calls(monkeypatch, graph, fuse):
    """Flow mode (one persistent all-reduce per step, layers picked up as the GEMM epilogues ring their
    device-side doorbells) against one fused call per layer: same weights bit for bit, no flag wait
    timed out.  On one rank this exercises exactly the producer -> consumer interlock: the persistent
    kernel is launched BEFORE the GEMMs that feed it."""
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    monkeypatch.setenv("AINIC_TIMEOUT_MS", "3000")
    monkeypatch.setenv("AINIC_MLP_GRAPH", graph)
    sizes = [256, 384, 264, 256, 128]
    cfg = MLPConfig(sizes=sizes, mb=256, fuse=fuse, lr=0.5, loss_weight=1.0, bucket_mb=0.0)
    g = torch.Generator().manual_seed(31)
    xs = [K.cast_bf16((torch.rand(cfg.mb, sizes[0], generator=g) - 0.5).cuda()) for _ in range(2)]
    ys = [torch.randint(0, sizes[-1], (cfg.mb,), generator=g, dtype=torch.int32).cuda() for _ in range(2)]
    out = {}
    for flow in ("0", "1"):
        monkeypatch.setenv("AINIC_FLOW", flow)
        tr = MLPTrainer(cfg)
        n0 = tr.comm.counters()["launches"]
        for i in range(5):
            tr.step(xs[i % 2], ys[i % 2])
        tr.wait()
        torch.cuda.synchronize()
        ctr = tr.check_health()
        out[flow] = ([tr.buffer("weight", l).cpu().clone() for l in range(cfg.nlayers)],
                     [tr.buffer("shadow", l).cpu().clone() for l in range(cfg.nlayers)], ctr["launches"] - n0)
        tr.close()
    assert out["0"][2] == 5 * cfg.nlayers and out["1"][2] == 5      # one launch per step in flow mode
    for a, b in zip(out["0"][0], out["1"][0]):
        assert torch.equal(a, b)
    for a, b in zip(out["0"][1], out["1"][1]):
        assert torch.equal(a, b)


def test_f32_mode_matches_fp32_oracle():
    """dtype="f32": fp32 activations and master weights straight into kind::tf32 tensor cores - the reference's
    all-fp32 MLP (libxsmm_dnn_fc_*_exec_f32).  Compared with the fp32 CPU oracle WITHOUT any bf16 rounding,
    at tf32 tolerance; the shadow-free update must leave the masters as the only copy."""
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    cfg = MLPConfig(sizes=[256, 384, 256, 128], mb=256, fuse=0, lr=0.5, loss_weight=1.0, dtype="f32")
    tr = MLPTrainer(cfg)
    g = torch.Generator().manual_seed(12)
    x = torch.rand(cfg.mb, cfg.sizes[0], generator=g) * 2 - 1
    y = torch.randint(0, cfg.sizes[-1], (cfg.mb,), generator=g, dtype=torch.int32)
    w0 = [tr.buffer("weight", l).cpu().clone() for l in range(cfg.nlayers)]
    xd, yd = x.cuda(), y.cuda()
    losses = []
    for _ in range(3):
        tr.step(xd, yd)
        tr.wait()
        torch.cuda.synchronize()
        losses.append(float(tr.loss_dev.item()) / cfg.mb)
    W = [w.clone() for w in w0]
    ref = []
    for _ in range(3):
        acts = [x]
        for l in range(cfg.nlayers):
            acts.append(acts[-1] @ W[l].t())
        loss, _, d = O.softmax_loss(acts[-1], y, cfg.loss_weight)
        ref.append(loss)
        for l in reversed(range(cfg.nlayers)):
            dw = d.t() @ acts[l]
            dn = d @ W[l]
            W[l] = O.sgd(W[l], dw, cfg.lr)
            d = dn
    for a, r in zip(losses, ref):
        assert abs(a - r) <= 5e-3 * abs(r) + 1e-3, (losses, ref)
    for l in range(cfg.nlayers):
        upd = (W[l] - w0[l]).norm().clamp_min(1e-12)
        assert ((tr.buffer("weight", l).cpu() - W[l]).norm() / upd).item() < 5e-3
    assert tr.buffer("act", 1).dtype == torch.float32
    tr.close()


def test_forward_mx_tracks_the_bf16_forward_and_the_weights():
    """Inference on the block-scaled fp8 tensor-core kind: same logits as the bf16 forward up to e4m3 rounding
    (one scale per 32 inputs), and the quantised weight copies follow the weights through a training step."""
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    cfg = MLPConfig(sizes=[256, 384, 256, 128], mb=256, fuse=0, lr=0.5, loss_weight=1.0)
    tr = MLPTrainer(cfg)
    g = torch.Generator().manual_seed(5)
    x = K.cast_bf16((torch.rand(cfg.mb, cfg.sizes[0], generator=g) - 0.5).cuda())
    y = torch.randint(0, cfg.sizes[-1], (cfg.mb,), generator=g, dtype=torch.int32).cuda()

    def rel(a, b):
        return ((a.float() - b.float()).norm() / b.float().norm()).item()

    ref0 = tr.forward(x).clone()
    mx0 = tr.forward_mx(x).clone()
    torch.cuda.synchronize()
    assert rel(mx0, ref0) < 0.08
    assert torch.equal(tr.forward_mx(x), mx0)            # deterministic, buffers reused
    for _ in range(3):
        tr.step(x, y)
    tr.wait()
    ref1 = tr.forward(x).clone()
    mx1 = tr.forward_mx(x).clone()
    torch.cuda.synchronize()
    assert rel(ref1, ref0) > 0.01                        # the steps did move the weights
    assert rel(mx1, ref1) < 0.08 and rel(mx1, ref1) < rel(mx0, ref1)
    tr.close()


def test_two_trainers_in_sequence_reuse_the_heap():
    """The symmetric heap is a real allocator now: closing a trainer returns its arenas, so a second trainer fits
    into a communicator sized for one (round 1's bump pointer would have been exhausted), lands on the same offsets
    and trains to the same bits; a staging area outgrown by a larger call is freed on a single rank."""
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    from fpga_ai_nic_b200.parallel.comm import Communicator
    cfg = MLPConfig(sizes=[512, 1024, 1024, 512], mb=256, fuse=0, lr=0.5, loss_weight=1.0)   # 2 M parameters: 22 MB of arenas
    comm = Communicator(cfg.heap_bytes(1))                                                    # + 8 MiB of slack
    base = comm.heap_stats()
    g = torch.Generator().manual_seed(9)
    x = K.cast_bf16((torch.rand(cfg.mb, cfg.sizes[0], generator=g) - 0.5).cuda())
    y = torch.randint(0, cfg.sizes[-1], (cfg.mb,), generator=g, dtype=torch.int32).cuda()
    outs, ptrs = [], []
    for _ in range(3):                      # three trainers, one after the other, in a heap sized for one
        tr = MLPTrainer(cfg, comm=comm)
        ptrs.append(tr.buffer("weight", 0).data_ptr())
        for _ in range(3):
            tr.step(x, y)
        tr.wait()
        torch.cuda.synchronize()
        tr.check_health()
        outs.append([tr.buffer("weight", l).cpu().clone() for l in range(cfg.nlayers)])
        tr.close()
    assert ptrs[0] == ptrs[1] == ptrs[2]
    for o in outs[1:]:
        for a, b in zip(outs[0], o):
            assert torch.equal(a, b)
    after = comm.heap_stats()
    assert after["in_use"] <= base["in_use"] + (1024 * 1024 // 512) * 6144 + (1 << 20)     # only the communicator's staging (largest layer) is left
    comm.close()
    # staging growth on one rank: the outgrown area goes straight back to the heap
    comm = Communicator(128 << 20)
    small = comm.empty(1 << 16)
    big = comm.empty(1 << 22)
    w = torch.zeros(1 << 22, device="cuda")
    s0 = comm.heap_stats()
    comm.allreduce_fused(small, w[: 1 << 16], 0.1, codec=K.CODEC_BFP)
    comm.allreduce_fused(big, w, 0.1, codec=K.CODEC_BFP)
    comm.allreduce_fused(big, w, 0.1, codec=K.CODEC_BFP)
    torch.cuda.synchronize()
    s1 = comm.heap_stats()
    frames = (1 << 22) // 512
    assert s1["in_use"] - s0["in_use"] <= frames * 6144 + (2 << 20)      # one staging area, not the sum of all sizes
    comm.free(big)
    comm.free(small)
    assert comm.heap_stats()["in_use"] == s1["in_use"] - (4 << 22) - (4 << 16)
    comm.close()
