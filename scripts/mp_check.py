"""Multi-process (one rank per GPU) correctness check, run under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 scripts/mp_check.py

1. fused all-reduce + SGD over IPC-mapped peer memory, every algorithm x codec, bit-exact against
   the CPU oracle (gradients of all ranks are gathered with NCCL only to build the expectation);
2. the NIC command interface on the GPU backend (10 requests, 8 done slots, in order);
3. the native DP trainer: replicas bit-identical after several steps, loss falls, and the weights
   match a CPU oracle run of the same data-parallel SGD within bf16 GEMM tolerance;
   3b. gradient buckets of several layers, the compressed GEMM epilogue and the GEMM -> reduce-scatter
   epilogue give the same bits as one call per layer on fp32 gradients;
4. start-up plumbing: parameter broadcast from rank 0, scatter of a global minibatch
   (the reference's MPI_Bcast / MPI_Scatter, sw/mlp_mpi_example_f32.cpp:453-470).
Prints one JSON line per check on rank 0 and exits non-zero on failure.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer  # noqa: E402
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402
from fpga_ai_nic_b200.ops import oracle as O  # noqa: E402
from fpga_ai_nic_b200.parallel.comm import Communicator  # noqa: E402
from fpga_ai_nic_b200.parallel.nic import Nic  # noqa: E402

rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
fails = []


def say(**kw):
    if rank == 0:
        print(json.dumps(kw), flush=True)


def gather(t: torch.Tensor) -> torch.Tensor:
    if world == 1:
        return t[None].cpu()
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return torch.stack(out).cpu()


def all_ok(ok: bool) -> bool:
    if world == 1:
        return ok
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


comm = Communicator(512 << 20)

# ------------------------------------------------------------------ 1. fused all-reduce + SGD
CASES = [("oneshot", "f32", 0), ("oneshot", "bfp", O.CLEAN), ("oneshot", "bfp", O.HW_EXACT), ("oneshot", "mxfp8", 0),
         ("oneshot", "bf16", 0), ("twoshot", "f32", 0), ("twoshot", "bfp", O.CLEAN), ("twoshot", "bfp", O.HW_EXACT),
         ("twoshot", "mxfp8", 0), ("twoshot", "bf16", 0), ("ring", "f32", 0), ("ring", "bfp", O.HW_EXACT)]
count_max = 1 << 20
g_heap = comm.empty(count_max)
w_heap = comm.empty(count_max)
sh = torch.empty(count_max, dtype=torch.bfloat16, device="cuda")
for count in (8192 * 5 + 777, 1 << 20):
    for it, (algo, codec, mode) in enumerate(CASES):
        gen = torch.Generator().manual_seed(100 * it + rank)
        g = torch.randn(count, generator=gen) * 1e-3
        w0 = torch.randn(count, generator=torch.Generator().manual_seed(7))
        gv, wv = g_heap[:count], w_heap[:count]
        gv.copy_(g)
        wv.copy_(w0)
        all_g = gather(gv)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        comm.allreduce_fused(gv, wv, 0.1, sh[:count], codec=K.CODEC_NAMES[codec], bfp_mode=mode, algo=K.ALGO_NAMES[algo])
        torch.cuda.synchronize()
        want = O.allreduce(all_g, "exact" if (codec == "f32" and algo != "ring") else algo, K.CODEC_NAMES[codec], mode)
        w_want = O.sgd(w0, want[rank], 0.1)
        ok = torch.equal(wv.cpu(), w_want) and torch.equal(sh[:count].cpu(), w_want.to(torch.bfloat16))
        ok = all_ok(ok)
        ctr = comm.counters()
        say(check="allreduce_fused", algo=algo, codec=codec, mode=mode, count=count, world=world, ok=ok,
            timeouts=ctr["timeouts"], last_us=ctr["last_ns"] / 1e3)
        if not ok or ctr["timeouts"]:
            fails.append(("allreduce", algo, codec, mode, count))
        if world > 1:
            dist.barrier()

# ------------------------------------------------------------------ 1a. gradient pre-compressed by the producer
# (what the weight-gradient GEMM epilogue does in the trainer): the kernel skips its own compression
count = 1 << 20
cbufs = [comm.empty(K.bfp_bytes(count), dtype=torch.uint8) for _ in range(2)]
for it, (algo, codec, mode) in enumerate([("oneshot", "bfp", O.CLEAN), ("twoshot", "bfp", O.CLEAN), ("twoshot", "mxfp8", 0),
                                          ("oneshot", "bfp", O.HW_EXACT)]):
    gen = torch.Generator().manual_seed(900 + 10 * it + rank)
    g = torch.randn(count, generator=gen) * 1e-3
    w0 = torch.randn(count, generator=torch.Generator().manual_seed(9))
    gv, wv = g_heap[:count], w_heap[:count]
    gv.copy_(g)
    wv.copy_(w0)
    cb = cbufs[it & 1]
    if codec == "bfp":
        K.bfp_compress(gv, mode, out=cb)
    else:
        K.mx_compress(gv, out=cb[:K.mx_bytes(count)])
    all_g = gather(gv)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    comm.allreduce_fused(gv, wv, 0.1, sh[:count], codec=K.CODEC_NAMES[codec], bfp_mode=mode, algo=K.ALGO_NAMES[algo],
                         grad_compressed=cb)
    torch.cuda.synchronize()
    want = O.allreduce(all_g, algo, K.CODEC_NAMES[codec], mode)
    ok = all_ok(torch.equal(wv.cpu(), O.sgd(w0, want[rank], 0.1)))
    say(check="allreduce_precompressed", algo=algo, codec=codec, mode=mode, world=world, ok=ok, timeouts=comm.counters()["timeouts"])
    if not ok:
        fails.append(("precompressed", algo, codec))
    if world > 1:
        dist.barrier()

# ------------------------------------------------------------------ 1b. NVLS (multimem) all-reduce + SGD
if world > 1:
    have = comm.enable_nvls(64 << 20)
    if not have:
        say(check="nvls", ok=True, available=False, reason=comm.nvls_error)
    else:
        count = 1 << 20
        gn = comm.nvls_empty(count)
        gen = torch.Generator().manual_seed(500 + rank)
        g = torch.randn(count, generator=gen) * 1e-3
        w0 = torch.randn(count, generator=torch.Generator().manual_seed(8))
        ok = True
        for rep in range(3):
            gn.copy_(g)
            wv = w_heap[:count]
            wv.copy_(w0)
            all_g = gather(gn)
            torch.cuda.synchronize()
            dist.barrier()
            comm.allreduce_nvls(gn, wv, 0.1, sh[:count])
            torch.cuda.synchronize()
            exact = all_g.double().sum(0).float()
            # the switch's summation order is not specified: compare with tolerance, replicas must agree
            w_want = O.sgd(w0, exact, 0.1)
            err = (wv.cpu() - w_want).abs().max().item()
            allw = gather(wv)
            ident = all(torch.equal(allw[r], allw[0]) for r in range(world))
            sums = gather(gn)
            ident &= all(torch.equal(sums[r], sums[0]) for r in range(world))
            ok &= err < 1e-6 and ident
        ok = all_ok(ok)
        ctr = comm.counters()
        say(check="nvls", ok=ok, available=True, max_err=err, identical=ident, timeouts=ctr["timeouts"], last_us=ctr["last_ns"] / 1e3)
        if not ok or ctr["timeouts"]:
            fails.append(("nvls",))
        # bf16 bucket: the switch returns bf16 sums; tolerance = a few bf16 roundings relative to the terms
        count = (1 << 20) - 512 * 3          # not a multiple of the 4-unit batch or of the shard size
        gb = comm.nvls_empty(count, torch.bfloat16)
        gen = torch.Generator().manual_seed(900 + rank)
        g16 = (torch.randn(count, generator=gen) * 1e-2).bfloat16()
        w0 = torch.randn(count, generator=torch.Generator().manual_seed(9))
        ok = True
        for rep in range(3):
            gb.copy_(g16)
            wv = w_heap[:count]
            wv.copy_(w0)
            all_g = gather(gb)
            torch.cuda.synchronize()
            dist.barrier()
            comm.allreduce_nvls(gb, wv, 0.1, sh[:count])
            torch.cuda.synchronize()
            exact = all_g.double().sum(0)
            sums = gather(gb)
            ident = all(torch.equal(sums[r], sums[0]) for r in range(world))
            # one bf16 rounding of the sum; measured against the magnitude of the terms, not of the sum (with 1 M
            # random elements some sums cancel almost completely and any relative-to-sum figure is meaningless)
            rel = ((sums[0].double() - exact).abs() / all_g.double().abs().sum(0).clamp_min(1e-30)).max().item()
            w_want = O.sgd(w0, sums[0].float().cpu(), 0.1)           # every rank applies the multicast bf16 sum
            err = (wv.cpu() - w_want).abs().max().item()
            allw = gather(wv)
            ident &= all(torch.equal(allw[r], allw[0]) for r in range(world))
            shd = sh[:count].cpu().float()
            # measured: 2^-7.04 at 2, 4 and 8 ranks - one full bf16 ulp: the in-switch result is truncated, not
            # rounded to nearest
            ok &= rel <= 2.0 ** -6.5 and err == 0.0 and ident and torch.equal(shd, wv.cpu().bfloat16().float())
        ok = all_ok(ok)
        ctr = comm.counters()
        say(check="nvls_bf16", ok=ok, sum_rel_err=rel, weight_err_vs_applied_sum=err, identical=ident, timeouts=ctr["timeouts"],
            last_us=ctr["last_ns"] / 1e3)
        if not ok or ctr["timeouts"]:
            fails.append(("nvls_bf16",))

# ------------------------------------------------------------------ 2. NIC command interface, GPU backend
nic = Nic("gpu", comm=comm, rank=rank, nranks=world)
count = 1 << 16
nic.setup(count, codec="bfp", bfp_mode=O.CLEAN, algo="twoshot", lr=0.5)
bufs = [comm.empty(count) for _ in range(10)]
wts = [comm.empty(count) for _ in range(10)]
for i in range(10):
    bufs[i].fill_(float(i + 1) * (rank + 1))
    wts[i].zero_()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
rqs = []
ok = True
for i in range(10):
    if i >= 8:
        nic.wait(rqs[i - 8], 30)          # at most 8 outstanding
    rqs.append(nic.all_reduce(bufs[i], wts[i], flags=int(i == 0)))
    ok &= rqs[-1].idx == i % 8
for i in range(10):
    nic.wait(rqs[i], 30)
torch.cuda.synchronize()
tot = sum(r + 1 for r in range(world))
for i in range(10):
    ok &= bool(torch.allclose(wts[i], torch.full_like(wts[i], -0.5 * (i + 1) * tot), rtol=2e-2))
rq = nic.iallreduce(bufs[0])
nic.wait(rq, 30)
torch.cuda.synchronize()
ok &= bool(torch.allclose(bufs[0], torch.full_like(bufs[0], 1.0 * tot), rtol=2e-2))
ok = all_ok(ok)
say(check="nic_gpu", ok=ok, latency_us=nic.latency_ns() / 1e3, stall_us=nic.stall_ns() / 1e3)
if not ok:
    fails.append(("nic",))
nic.close()

# ------------------------------------------------------------------ 3. native DP trainer
for codec in ("f32", "bfp"):
    cfg = MLPConfig(sizes=[256, 512, 384, 128], mb=256, fuse=0, lr=0.5, loss_weight=1.0, codec=codec)
    tr = MLPTrainer(cfg, comm=comm)
    gen = torch.Generator().manual_seed(4321 + rank)
    x = torch.rand(cfg.mb, cfg.sizes[0], generator=gen) * 2 - 1
    y = torch.randint(0, cfg.sizes[-1], (cfg.mb,), generator=gen, dtype=torch.int32)
    xb, yd = K.cast_bf16(x.cuda()), y.cuda()
    w_init = [tr.buffer("weight", l).cpu().clone() for l in range(cfg.nlayers)]
    losses = []
    steps = 4
    for _ in range(steps):
        tr.step(xb, yd)
        tr.wait()
        torch.cuda.synchronize()
        losses.append(float(tr.loss_dev.item()) / cfg.mb)
    ident = True
    for l in range(cfg.nlayers):
        allw = gather(tr.buffer("weight", l).reshape(-1))
        ident &= all(torch.equal(allw[r], allw[0]) for r in range(world))
    # CPU oracle of the same data-parallel SGD (fp32 everywhere)
    xs, ys = gather(x.cuda()), gather(y.cuda())
    W = [w.clone() for w in w_init]
    for _ in range(steps):
        grads = [torch.zeros_like(w) for w in W]
        for r in range(world):
            acts = [xs[r].bfloat16().float()]
            for l in range(cfg.nlayers):
                acts.append(O.fc_fwd(acts[-1], W[l].bfloat16().float())[0].bfloat16().float())
            _, _, d = O.softmax_loss(acts[-1], ys[r], cfg.loss_weight)
            d = d.bfloat16().float()
            for l in reversed(range(cfg.nlayers)):
                dx, dw, _ = O.fc_bwd(acts[l], W[l].bfloat16().float(), d, need_dx=l > 0)
                grads[l] += dw
                if l > 0:
                    d = dx.bfloat16().float()
        W = [O.sgd(W[l], grads[l], cfg.lr) for l in range(cfg.nlayers)]
    rel = max(((tr.buffer("weight", l).cpu() - W[l]).norm() / (W[l] - w_init[l]).norm().clamp_min(1e-12)).item()
              for l in range(cfg.nlayers))
    ok = ident and losses[-1] < losses[0] and rel < (0.02 if codec == "f32" else 0.08)
    ok = all_ok(ok)
    say(check="mlp_trainer", codec=codec, world=world, ok=ok, identical_replicas=ident, losses=losses,
        update_rel_err_vs_oracle=rel, timeouts=comm.counters()["timeouts"])
    if not ok:
        fails.append(("mlp", codec))
    tr.close()
    # 3b. gradient bucketing, the compressed GEMM epilogue (1) and the GEMM -> reduce-scatter epilogue (2)
    # must give the same bits as one call per layer on fp32 gradients: the sum order and the codec
    # blocks do not depend on which rank owns a frame or on who compressed it.  The algorithm is pinned
    # (auto would pick one-shot for small buckets and two-shot for large ones, which round differently).
    algo = "twoshot" if world > 1 else "auto"

    def run_variant(bucket_mb, fusec):
        os.environ["AINIC_FUSE_COMPRESS"] = fusec
        tb = MLPTrainer(MLPConfig(**{**cfg.__dict__, "bucket_mb": bucket_mb, "algo": algo}), comm=comm)
        os.environ.pop("AINIC_FUSE_COMPRESS", None)
        for _ in range(steps):
            tb.step(xb, yd)
        tb.wait()
        torch.cuda.synchronize()
        out = [tb.buffer("weight", l).cpu().clone() for l in range(cfg.nlayers)]
        tb.close()
        return out

    base = run_variant(0.0, "0")
    variants = [(1000.0, "0"), (0.6, "0")]
    if codec == "bfp" and world > 1:
        variants += [(0.0, "1"), (1000.0, "1"), (0.0, "2"), (0.6, "2"), (1000.0, "2")]
    for bucket_mb, fusec in variants:
        got = run_variant(bucket_mb, fusec)
        same = all_ok(all(torch.equal(got[l], base[l]) for l in range(cfg.nlayers)))
        say(check="mlp_bucketed", codec=codec, bucket_mb=bucket_mb, fuse_compress=fusec, ok=same,
            timeouts=comm.counters()["timeouts"])
        if not same:
            fails.append(("mlp_bucketed", codec, bucket_mb, fusec))

# ------------------------------------------------------------------ 4. start-up plumbing: Bcast of parameters, Scatter of a batch
cfg = MLPConfig(sizes=[256, 256, 128], mb=64, codec="f32")
tr = MLPTrainer(cfg, comm=comm)
ref_w = [tr.buffer("weight", l).clone() for l in range(cfg.nlayers)]
if rank != 0:
    for l in range(cfg.nlayers):
        tr.buffer("weight", l).add_(1.0 + rank)          # diverge on purpose
tr.broadcast_parameters(0)
ok = all(torch.equal(tr.buffer("weight", l), ref_w[l]) for l in range(cfg.nlayers))
ok &= all(torch.equal(tr.buffer("shadow", l), ref_w[l].bfloat16()) for l in range(cfg.nlayers))
gx = torch.arange(world * cfg.mb * 256, dtype=torch.float32).reshape(world * cfg.mb, 256) % 251 if rank == 0 else None
gy = (torch.arange(world * cfg.mb, dtype=torch.int32) % 128) if rank == 0 else None
xs, ys = tr.scatter_batch(gx, gy, 0)
want_x = (torch.arange(world * cfg.mb * 256, dtype=torch.float32).reshape(world, cfg.mb, 256) % 251)[rank].bfloat16()
want_y = (torch.arange(world * cfg.mb, dtype=torch.int32) % 128).reshape(world, cfg.mb)[rank]
ok &= torch.equal(xs.cpu(), want_x) and torch.equal(ys.cpu(), want_y)
ok = all_ok(bool(ok))
say(check="bcast_scatter", ok=ok)
if not ok:
    fails.append(("bcast_scatter",))
tr.close()

if world > 1:
    dist.barrier()
    dist.destroy_process_group()
say(check="summary", world=world, failures=fails)
sys.exit(1 if fails else 0)
