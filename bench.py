#!/usr/bin/env python
"""Headline benchmark: MLP training samples/s (BASELINE.json: "MLP train samples/sec @1/2/4/8 B200").

Config = the reference's own run (sw/run.sh:16): 10 fully-connected layers 2048x2048, 1792 samples
per rank (weak scaling), fuse 0, SGD with lr 0.1 folded into the gradient all-reduce.  bf16 tensor
core GEMMs with fp32 accumulation and fp32 master weights; synthetic data, random-init weights.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 5
    python bench.py --impl ddp ...        # PyTorch DDP (NCCL + cuBLAS + separate SGD) baseline
    python bench.py --impl reference ...  # the unmodified FPGA reference: unavailable

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# Load every kernel when the CUDA context is created: a persistent kernel that waits for a producer must
# never be the reason that producer's module load stalls (see fpga_ai_nic_b200/__init__.py).
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "ddp", "reference"])
    ap.add_argument("--layers", type=int, default=10)
    ap.add_argument("--width", type=int, default=2048)
    ap.add_argument("--mb", type=int, default=1792, help="samples per rank (reference: 1792)")
    ap.add_argument("--codec", default="auto", choices=["auto", "f32", "bfp", "mxfp8", "bf16"])
    ap.add_argument("--algo", default="auto", choices=["auto", "oneshot", "twoshot", "ring"])
    ap.add_argument("--fuse", type=int, default=0)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "f32"],
                    help="bf16: bf16 activations / weight shadows (headline); f32: the reference's all-fp32 data flow on tf32 tensor cores")
    ap.add_argument("--bucket-mb", type=float, default=None,
                    help="fp32 MiB of gradients per fused all-reduce (0 = per layer; default: 48 at 2 ranks, 80 from 3)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-baseline", action="store_true",
                    help="--impl ours: skip the PyTorch DDP arms that are otherwise measured in the same invocation")
    ap.add_argument("--no-check", action="store_true", help="skip the replica / update-error checks at N > 1")
    ap.add_argument("--ddp-bucket-mb", type=float, default=34.0, help="bucket_cap_mb of the hardened DDP arm")
    ap.add_argument("--ddp-mode", default="naive", choices=["naive", "hardened", "hardened-eager"],
                    help="--impl ddp: which DDP arm (hardened-eager = hardened without the CUDA-graph capture)")
    return ap.parse_args()


def reference_arm(args):
    """The reference is SystemVerilog for an Arria-10 PAC plus one C++ file that needs OPAE, Intel MPI
    and libxsmm_dnn; it has no Python packaging, so `pip install /root/reference` cannot produce
    baseline/_ref (outcome recorded in DESIGN.md)."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    why = ("reference is FPGA RTL (hw/*.sv) + an OPAE/MPI/libxsmm C++ app with no setup.py/pyproject; "
           "pip install into baseline/_ref fails and no FPGA/OPAE/MPI exists on this box")
    if os.path.isdir(ref_dir) and os.listdir(ref_dir):
        why = "baseline/_ref exists but holds no runnable entry point: " + why
    print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


def maybe_respawn(args):
    """`python bench.py --gpus N` without torchrun: re-launch ourselves under torch.distributed.run."""
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        import subprocess
        port = 29500 + (os.getpid() % 2000)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))


def dist_setup(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, local, world


def max_over_ranks(x: float, world: int) -> float:
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world: int):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


LAST_STEP_MS: list = []   # per-step device times of the most recent timed_loop on this rank (diagnostic)


def timed_loop(step_fn, steps: int, warmup: int, world: int, local: int, finalize=None):
    """W untimed steps, then exactly K steps between barrier+synchronize, CUDA events, max over ranks.
    `finalize` (optional) runs after the last step but before the closing event (pipelined read-backs)."""
    import torch
    from fpga_ai_nic_b200.utils.clocks import ClockSampler
    # the sampler thread starts (NVML init, first queries) before the warm-up, not inside the timed region
    sampler = ClockSampler(local, period_ms=5.0).start()
    for i in range(warmup):
        step_fn(i)
    if finalize is not None:
        finalize()
    barrier(world)
    sampler.mark()
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    marks[0].record()
    for i in range(steps):
        step_fn(warmup + i)
        if i + 1 < steps:
            marks[i + 1].record()
    if finalize is not None:
        finalize()
    marks[steps].record()
    barrier(world)
    clocks = sampler.stop()
    ms = max_over_ranks(marks[0].elapsed_time(marks[steps]), world)
    LAST_STEP_MS[:] = [round(marks[i].elapsed_time(marks[i + 1]), 4) for i in range(steps)]
    return ms, clocks


def run_ours(args, rank, local, world):
    import torch
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    cfg = MLPConfig.reference(args.layers, args.width, args.mb)
    cfg.fuse, cfg.codec, cfg.algo, cfg.bucket_mb, cfg.dtype = args.fuse, args.codec, args.algo, args.bucket_mb, args.dtype
    f32 = args.dtype == "f32"
    tr = MLPTrainer(cfg)
    dev = tr.device
    g = torch.Generator().manual_seed(1234 + rank)
    nbuf = 4  # rotating synthetic batches, fp32 on the host like the reference's activations
    host_x = [(torch.rand(cfg.mb, cfg.sizes[0], generator=g) * 0.1 - 0.05).pin_memory() for _ in range(nbuf)]
    host_y = [torch.randint(0, cfg.sizes[-1], (cfg.mb,), generator=g, dtype=torch.int32).pin_memory() for _ in range(nbuf)]
    dev_x = [h.to(dev) if f32 else K.cast_bf16(h.to(dev)) for h in host_x]
    dev_y = [h.to(dev) for h in host_y]

    # ---- device-resident inputs: the kernel-level number
    def step_dev(i):
        tr.step(dev_x[i % nbuf], dev_y[i % nbuf], want_loss=True)

    l0 = tr.launches
    ms, clocks = timed_loop(lambda i: step_dev(i), args.steps, args.warmup, world, local)
    step_ms = list(LAST_STEP_MS)
    tr.wait()
    torch.cuda.synchronize()
    launches = (tr.launches - l0) * args.steps // (args.steps + args.warmup)
    loss = float(tr.loss_dev.item()) / cfg.mb

    # ---- end to end through the public API: pinned host -> device every step, loss read back every step
    e2e = None
    if not args.no_e2e:
        copy_stream = torch.cuda.Stream()
        loss_stream = torch.cuda.Stream()
        probe_mode = os.environ.get("AINIC_BENCH_PROBE", "")

        def run_e2e(hx, steps, warmup):
            """hx: pinned host batches (fp32 like the reference's host activations, or bf16 as a loader that
            already emits the compute dtype would provide)."""
            narrow = hx[0].dtype == torch.bfloat16 or f32      # staged batch is already in the compute dtype
            stage_in = [torch.empty(cfg.mb, cfg.sizes[0], dtype=hx[0].dtype, device=dev) for _ in range(2)]
            stage_bf16 = stage_in if narrow else [torch.empty(cfg.mb, cfg.sizes[0], dtype=torch.bfloat16, device=dev) for _ in range(2)]
            stage_y = [torch.empty(cfg.mb, dtype=torch.int32, device=dev) for _ in range(2)]
            ready = [torch.cuda.Event() for _ in range(2)]
            consumed = [torch.cuda.Event() for _ in range(2)]
            step_done = [torch.cuda.Event() for _ in range(2)]
            loss_dev2 = [torch.zeros(1, dtype=torch.float32, device=dev) for _ in range(2)]
            host_loss = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
            loss_ready = [torch.cuda.Event() for _ in range(2)]
            losses = []

            def upload(i):
                s = i % 2
                with torch.cuda.stream(copy_stream):
                    copy_stream.wait_event(consumed[s])
                    if probe_mode == "noh2d" and i >= 2:      # diagnostic only (AINIC_BENCH_PROBE): keeps the pipeline, drops the copy
                        ready[s].record(copy_stream)
                        return
                    stage_in[s].copy_(hx[i % nbuf], non_blocking=True)
                    stage_y[s].copy_(host_y[i % nbuf], non_blocking=True)
                    if not narrow:
                        K.cast_bf16(stage_in[s], out=stage_bf16[s])
                    ready[s].record(copy_stream)

            state = {"primed": False, "pending": None}

            def read_pending():
                if state["pending"] is not None:
                    p = state["pending"]
                    loss_ready[p].synchronize()
                    losses.append(float(host_loss[p].item()) / cfg.mb)
                    state["pending"] = None

            def step_e2e(i):
                main = torch.cuda.current_stream()
                if not state["primed"]:
                    upload(i)
                    state["primed"] = True
                s = i % 2
                main.wait_event(ready[s])
                upload(i + 1)                      # prefetch the next batch while this step computes
                main.wait_event(loss_ready[s])     # loss slot s was read out two steps ago
                tr.step(stage_bf16[s], stage_y[s], want_loss=True, loss_out=loss_dev2[s])
                consumed[s].record(main)
                step_done[s].record(main)
                # device -> host copy of the loss on a side stream: a 4-byte DMA costs ~10 us of stream time,
                # which the main stream would otherwise spend idle before the next forward
                with torch.cuda.stream(loss_stream):
                    loss_stream.wait_event(step_done[s])
                    host_loss[s].copy_(loss_dev2[s], non_blocking=True)
                    loss_ready[s].record(loss_stream)
                # read the PREVIOUS step's loss on the host: the CPU stays one step ahead of the GPU
                # instead of draining the pipeline every iteration
                if probe_mode != "noloss":
                    read_pending()
                state["pending"] = s

            ms_e, _ = timed_loop(step_e2e, steps, warmup, world, local, finalize=read_pending)
            return ms_e, losses

        ms_e2e, losses = run_e2e(host_x, args.steps, max(3, args.warmup))
        in_bytes = cfg.mb * cfg.sizes[0] * 4 + cfg.mb * 4
        # the host -> device copy alone (same buffers, nothing else running): the floor of an fp32-input step
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        probe = torch.empty(cfg.mb, cfg.sizes[0], device=dev)
        probe.copy_(host_x[0], non_blocking=True)
        torch.cuda.synchronize()
        barrier(world)
        t0.record()
        for i in range(20):
            probe.copy_(host_x[i % nbuf], non_blocking=True)
        t1.record()
        torch.cuda.synchronize()
        h2d_ms = max_over_ranks(t0.elapsed_time(t1) / 20, world)
        e2e = {"value": cfg.mb * world * args.steps / (ms_e2e / 1e3), "unit": "samples/s",
               "ms_per_step": ms_e2e / args.steps,
               "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": 4,
               "h2d_copy_alone_ms": h2d_ms, "h2d_copy_alone_gbps": in_bytes / h2d_ms / 1e6,
               "numa_bind": getattr(tr.comm, "numa", None),
               "note": "fp32 inputs from pinned host memory every step (prefetched one step ahead on a copy stream), "
                       "cast to bf16 on device; every step's loss is copied to pinned host memory and read by the "
                       "host one step later (the last one before the closing event)"}
        loss = losses[-1]
        # same loop with the host batches already in the compute dtype (half the bytes over PCIe)
        if not f32:
            host_xb = [h.bfloat16().pin_memory() for h in host_x]
            ms_b, losses_b = run_e2e(host_xb, args.steps, max(3, args.warmup))
            e2e["bf16_host_inputs"] = {"value": cfg.mb * world * args.steps / (ms_b / 1e3), "ms_per_step": ms_b / args.steps,
                                       "h2d_bytes_per_step": cfg.mb * cfg.sizes[0] * 2 + cfg.mb * 4, "loss": losses_b[-1]}
    ctr = tr.comm.counters()
    value = cfg.mb * world * args.steps / (ms / 1e3)
    gflops = cfg.flops_per_step(cfg.mb * world) / (ms / args.steps / 1e3) / 1e9
    out = {
        "metric": "mlp_train_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "tf32 (fp32 storage, kind::tf32 MMA)" if f32 else "bf16",
        "data": "synthetic", "impl": "ours",
        "config": {"model": f"mlp_{cfg.nlayers}x{args.width}x{args.width}", "global_batch": cfg.mb * world,
                   "seq_len": 1, "parallelism": f"dp{world}", "per_rank_batch": cfg.mb, "fuse": cfg.fuse,
                   "optimizer": "sgd lr=0.1 fused into the gradient all-reduce", "master_weights": "fp32",
                   "grad_codec": {0: "f32", 1: "bfp16x8", 2: "mxfp8", 3: "bf16"}[tr.codec], "algo": args.algo,
                   "grad_bucket_mib": tr.bucket_mb,
                   "l2": "no flush: per-step working set (~0.6 GB of weights, gradients, shadows, activations) exceeds the 126 MB L2"},
        "clocks": {"sm_mhz": clocks.get("sm_mhz"), "sm_max_mhz": clocks.get("sm_max_mhz"),
                   "reasons": clocks.get("reasons", []), "power_w_max": clocks.get("power_w_max")},
        "gpu_launches": int(launches), "e2e": e2e, "loss": loss,
        "gflops_reference_formula": gflops, "comm_counters": ctr, "step_ms": step_ms,
    }
    out["config"]["flow_mode"] = os.environ.get("AINIC_FLOW", "0 (default: one fused all-reduce call per gradient bucket)")
    if world > 1 and not args.no_check:
        try:
            out["correctness"] = check_replicas_and_update(args, tr, cfg, world, rank)
        except Exception as ex:   # never lose the bench line to the check
            out["correctness"] = {"error": f"{type(ex).__name__}: {str(ex)[:200]}"}
    tr.close(collective=False)   # the communicator (and with it the whole heap) goes away on the next line
    tr.comm.close()
    del dev_x, host_x
    torch.cuda.empty_cache()
    return out


def _pinned_batches(args, rank, nbuf=4):
    import torch
    g = torch.Generator().manual_seed(1234 + rank)
    xs = [(torch.rand(args.mb, args.width, generator=g) * 0.1 - 0.05).pin_memory() for _ in range(nbuf)]
    ys = [torch.randint(0, args.width, (args.mb,), generator=g, dtype=torch.int32).pin_memory() for _ in range(nbuf)]
    return xs, ys


def run_ddp(args, rank, local, world, mode="naive"):
    """PyTorch DDP baselines of the same MLP, same sizes, same data shape (the software path the reference's
    NIC replaces: MPI_Iallreduce + opt_exec, sw/mlp_mpi_example_f32.cpp:590-680, on this box NCCL + cuBLAS).

    naive     what a user writes first: fp32 parameters, bf16 autocast (re-casts every weight each step),
              fp32 cross-entropy, torch.optim.SGD, DDP default-ish buckets, eager launches.
    hardened  bf16 parameters + fp32 master copies updated with fused foreach ops, bf16 gradient
              all-reduce (half the wire bytes), static_graph DDP with gradient-as-bucket-view and the
              requested bucket size, and the whole step replayed as ONE CUDA graph when capture with NCCL
              succeeds (falls back to eager otherwise; the JSON says which ran).
    Both report a device-timed number on device-resident inputs and an end-to-end number with the inputs
    copied from pinned host memory every step (prefetched one step ahead) and the loss read back."""
    import torch
    import torch.nn as nn
    import torch.nn.functional as F
    dev = torch.device("cuda", local)
    if world > 1:   # same host-side placement as the product arm (Communicator binds its rank the same way)
        from fpga_ai_nic_b200.utils.numa import bind_to_gpu
        bind_to_gpu(local)
    torch.manual_seed(1)
    hardened = mode in ("hardened", "hardened-eager")
    try_graph = mode == "hardened"
    pdt = torch.bfloat16 if hardened else torch.float32
    layers = [nn.Linear(args.width, args.width, bias=(args.fuse in (1, 3))) for _ in range(args.layers)]
    seq = []
    for l in layers:
        seq.append(l)
        if args.fuse >= 2:
            seq.append(nn.ReLU())
    net = nn.Sequential(*seq).to(dev).to(pdt)
    model = net
    if world > 1:
        from torch.nn.parallel import DistributedDataParallel as DDP
        if hardened:
            model = DDP(net, device_ids=[local], bucket_cap_mb=args.ddp_bucket_mb, gradient_as_bucket_view=True,
                        static_graph=True)
        else:
            model = DDP(net, device_ids=[local], bucket_cap_mb=17, gradient_as_bucket_view=True)
    lr = 0.1 * 0.1
    params = [p for p in net.parameters()]
    if hardened:
        masters = [p.detach().float().clone() for p in params]
    else:
        opt = torch.optim.SGD(params, lr=lr)
    host_x, host_y = _pinned_batches(args, rank)
    xs = [h.to(dev).to(pdt) if hardened else h.to(dev) for h in host_x]
    ys = [h.to(dev).long() for h in host_y]
    static_x = torch.zeros_like(xs[0])
    static_y = torch.zeros_like(ys[0])
    loss_buf = torch.zeros((), device=dev)

    def train_step(x, y):
        if hardened:
            out = model(x)
            loss = F.cross_entropy(out.float(), y)
            for p in params:
                p.grad = None
            loss.backward()
            with torch.no_grad():
                grads = [p.grad for p in params]
                torch._foreach_add_(masters, [g.float() for g in grads], alpha=-lr)
                torch._foreach_copy_(params, masters)
        else:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = model(x)
            loss = F.cross_entropy(out.float(), y)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
        loss_buf.copy_(loss.detach())

    graph = None
    graphed = False
    if try_graph:
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(11):                      # DDP needs its warm-up iterations before a capture
                    train_step(static_x, static_y)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            # thread-local capture mode: the NCCL watchdog thread keeps querying events of earlier collectives,
            # which a global-mode capture treats as an illegal call and invalidates itself on (seen at 8 ranks)
            with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                train_step(static_x, static_y)
            graph.replay()
            torch.cuda.synchronize()
            graphed = True
        except Exception as ex:      # capture with NCCL is not guaranteed everywhere: eager fallback, recorded below
            graph = None
            graph_error = f"{type(ex).__name__}: {str(ex)[:160]}"
            torch.cuda.synchronize()
    ok = torch.tensor([1 if (graphed or not try_graph) else 0], device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)   # all ranks replay, or none does
    if try_graph and int(ok.item()) == 0:
        # a failed capture leaves DDP / NCCL in an undefined state (measured: 5.5 ms/step afterwards): say so and
        # stop; the caller re-runs this arm as "hardened-eager" in a fresh process
        if rank == 0:
            print(json.dumps({"impl": "ddp-hardened", "capture_failed": locals().get("graph_error", "a peer rank could not capture")}), flush=True)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        sys.exit(3)

    def step_dev(i):
        if graph is not None:
            static_x.copy_(xs[i % 4], non_blocking=True)
            static_y.copy_(ys[i % 4], non_blocking=True)
            graph.replay()
        else:
            train_step(xs[i % 4], ys[i % 4])

    ms, clocks = timed_loop(step_dev, args.steps, args.warmup, world, local)
    step_ms = list(LAST_STEP_MS)
    loss = float(loss_buf.item())

    # ---- end to end: pinned host -> device every step (prefetched on a copy stream), loss read one step later
    e2e = None
    if not args.no_e2e:
        copy_stream = torch.cuda.Stream()
        stage_x = [torch.empty(args.mb, args.width, device=dev) for _ in range(2)]
        stage_y = [torch.empty(args.mb, dtype=torch.int32, device=dev) for _ in range(2)]
        ready = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]
        host_loss = [torch.zeros(()).pin_memory() for _ in range(2)]
        loss_ready = [torch.cuda.Event() for _ in range(2)]
        st = {"primed": False, "pending": None}
        losses = []

        def upload(i):
            sl = i % 2
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[sl])
                stage_x[sl].copy_(host_x[i % 4], non_blocking=True)
                stage_y[sl].copy_(host_y[i % 4], non_blocking=True)
                ready[sl].record(copy_stream)

        def read_pending():
            if st["pending"] is not None:
                loss_ready[st["pending"]].synchronize()
                losses.append(float(host_loss[st["pending"]].item()))
                st["pending"] = None

        def step_e2e(i):
            main = torch.cuda.current_stream()
            if not st["primed"]:
                upload(i)
                st["primed"] = True
            sl = i % 2
            main.wait_event(ready[sl])
            upload(i + 1)
            if graph is not None:
                static_x.copy_(stage_x[sl])          # fp32 -> bf16 cast on device, like the product arm
                static_y.copy_(stage_y[sl])
                consumed[sl].record(main)
                graph.replay()
            else:
                x = stage_x[sl].to(pdt) if hardened else stage_x[sl]
                y = stage_y[sl].long()
                train_step(x, y)
                consumed[sl].record(main)
            host_loss[sl].copy_(loss_buf, non_blocking=True)
            loss_ready[sl].record(main)
            read_pending()
            st["pending"] = sl

        ms_e, _ = timed_loop(step_e2e, args.steps, max(3, args.warmup), world, local, finalize=read_pending)
        e2e = {"value": args.mb * world * args.steps / (ms_e / 1e3), "unit": "samples/s", "ms_per_step": ms_e / args.steps,
               "h2d_bytes_per_step": args.mb * args.width * 4 + args.mb * 4, "d2h_bytes_per_step": 4}
    value = args.mb * world * args.steps / (ms / 1e3)
    stack = ("torch DDP hardened: bf16 parameters + fp32 masters (foreach SGD), bf16 NCCL all-reduce, static_graph, "
             f"bucket_cap_mb={args.ddp_bucket_mb}, " + ("whole step replayed as one CUDA graph" if graphed else "eager launches")
             if hardened else "torch DDP naive: cuBLAS bf16 autocast + NCCL fp32 all-reduce (bucket_cap_mb=17) + torch.optim.SGD, eager")
    out = {
        "metric": "mlp_train_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "ddp-" + mode,
        "config": {"model": f"mlp_{args.layers}x{args.width}x{args.width}", "global_batch": args.mb * world,
                   "seq_len": 1, "parallelism": f"dp{world}", "per_rank_batch": args.mb, "stack": stack,
                   "cuda_graph": graphed},
        "clocks": {"sm_mhz": clocks.get("sm_mhz"), "sm_max_mhz": clocks.get("sm_max_mhz"), "reasons": clocks.get("reasons", [])},
        "gpu_launches": 0, "loss": loss, "e2e": e2e, "step_ms": step_ms,
    }
    if mode == "hardened-eager":
        out["config"]["cuda_graph_error"] = "capture skipped (an attempt in a separate process failed)"
    del model, net, params
    torch.cuda.empty_cache()
    return out


def measure_baselines(args, ours, rank, world):
    """The metric is "samples/sec vs NCCL DDP": measure that denominator in the same invocation - same
    box, same N, same steps / warm-up - right after the product arm.  Each DDP arm runs in a child
    process per rank (own rendezvous port), so nothing it does (a failed graph capture with NCCL, an
    out-of-memory) can take the product arm's line down with it.  vs_baseline = ours / the faster arm."""
    import subprocess
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    base = {}
    port0 = int(os.environ.get("MASTER_PORT", "29500"))
    modes = ["naive", "hardened"]
    idx = -1
    while modes:
        mode = modes.pop(0)
        idx += 1
        env = dict(os.environ)
        env["MASTER_ADDR"] = "127.0.0.1"
        env["MASTER_PORT"] = str(port0 + 101 + idx)
        env["TORCHELASTIC_USE_AGENT_STORE"] = "False"      # the child group's rank 0 hosts its own store
        cmd = [sys.executable, os.path.abspath(__file__), "--impl", "ddp", "--ddp-mode", mode, "--gpus", str(args.gpus),
               "--steps", str(args.steps), "--warmup", str(args.warmup), "--layers", str(args.layers), "--width", str(args.width),
               "--mb", str(args.mb), "--fuse", str(args.fuse), "--ddp-bucket-mb", str(args.ddp_bucket_mb)]
        if args.no_e2e:
            cmd.append("--no-e2e")
        try:
            p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=420)
            line = [l for l in p.stdout.splitlines() if l.startswith("{")]
            if rank == 0:
                if p.returncode == 0 and line:
                    d = json.loads(line[-1])
                    base[mode] = {"samples_per_s": d["value"], "ms_per_step": d["ms_per_step"],
                                  "e2e_samples_per_s": d["e2e"]["value"] if d.get("e2e") else None,
                                  "e2e_ms_per_step": d["e2e"]["ms_per_step"] if d.get("e2e") else None,
                                  "config": d["config"]["stack"], "cuda_graph": d["config"].get("cuda_graph"),
                                  "cuda_graph_error": d["config"].get("cuda_graph_error"), "loss": d["loss"],
                                  "clocks": d.get("clocks")}
                else:
                    base[mode] = {"error": f"rc={p.returncode}: {p.stderr[-300:]}"}
            if mode == "hardened" and p.returncode == 3:
                modes.append("hardened-eager")       # every rank sees the same exit code: all of them retry
        except Exception as ex:
            base[mode] = {"error": f"{type(ex).__name__}: {str(ex)[:200]}"}
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
    if rank != 0:
        return {}
    ok = {m: b for m, b in base.items() if "samples_per_s" in b}
    best = max((b["samples_per_s"] for b in ok.values()), default=None)
    best_e2e = max((b["e2e_samples_per_s"] for b in ok.values() if b.get("e2e_samples_per_s")), default=None)
    res = {"baseline": {"what": "PyTorch DDP (NCCL + cuBLAS) of the same MLP, measured in this invocation right after the "
                                "product arm (child processes, same GPUs); vs_baseline = value / the faster of the two arms",
                        "ddp": base,
                        "ratio_vs_naive": ours["value"] / ok["naive"]["samples_per_s"] if "naive" in ok else None,
                        "ratio_vs_hardened": (ours["value"] / max(ok[m]["samples_per_s"] for m in ("hardened", "hardened-eager") if m in ok))
                                             if ("hardened" in ok or "hardened-eager" in ok) else None,
                        "e2e_ratio_vs_best": (ours["e2e"]["value"] / best_e2e) if (ours.get("e2e") and best_e2e) else None}}
    if best:
        res["vs_baseline"] = ours["value"] / best
    return res


def check_replicas_and_update(args, tr, cfg, world, rank):
    """Driver-visible correctness at N > 1 (the reference's cluster run is judged by completion only,
    sw/run.sh:16).  (1) every rank's fp32 master weights are bit-identical after the timed loops;
    (2) one more step on a fixed batch with the product's gradient codec vs the same step with an
    uncompressed fp32 all-reduce: relative L2 error of the applied update, per layer (max reported)."""
    import torch
    import torch.distributed as dist
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    tr.wait()
    torch.cuda.synchronize()
    dev = tr.device
    def checksum(w):   # two order-sensitive 64-bit sums over the raw bits
        flat = w.view(torch.int32).flatten().to(torch.int64)
        k = torch.arange(flat.numel(), device=flat.device, dtype=torch.int64) % 1021 + 1
        return torch.stack([flat.sum(), (flat * k).sum()])
    sig = torch.cat([checksum(tr.buffer("weight", l)) for l in range(cfg.nlayers)])
    allsig = [torch.zeros_like(sig) for _ in range(world)]
    dist.all_gather(allsig, sig)
    identical = all(torch.equal(allsig[0], t) for t in allsig)
    # (2) same weights, same batch, fp32 all-reduce trainer as the yardstick
    cfg2 = MLPConfig.reference(args.layers, args.width, args.mb)
    cfg2.fuse, cfg2.codec, cfg2.algo, cfg2.bucket_mb = args.fuse, "f32", "auto", args.bucket_mb
    ref = MLPTrainer(cfg2)
    g = torch.Generator().manual_seed(99 + rank)
    x = K.cast_bf16((torch.rand(cfg.mb, cfg.sizes[0], generator=g) * 0.1 - 0.05).to(dev))
    y = torch.randint(0, cfg.sizes[-1], (cfg.mb,), generator=g, dtype=torch.int32).to(dev)
    w0 = [tr.buffer("weight", l).clone() for l in range(cfg.nlayers)]
    for l in range(cfg.nlayers):
        ref.buffer("weight", l).copy_(w0[l])
    ref.sync_shadows()
    for t in (tr, ref):
        t.step(x, y)
        t.wait()
    torch.cuda.synchronize()
    errs = []
    for l in range(cfg.nlayers):
        exact = ref.buffer("weight", l) - w0[l]
        ours = tr.buffer("weight", l) - w0[l]
        errs.append(float(((ours - exact).norm() / exact.norm().clamp_min(1e-30)).item()))
    timeouts = ref.comm.counters()["timeouts"]
    ref.close(collective=False)
    ref.comm.close()
    t = torch.tensor([max(errs)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return {"replicas_identical": bool(identical), "update_rel_err_vs_fp32_allreduce": float(t.item()),
            "update_rel_err_per_layer": [round(e, 5) for e in errs], "check_timeouts": int(timeouts)}


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    maybe_respawn(args)
    rank, local, world = dist_setup(args)
    if world != args.gpus and rank == 0:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}", file=sys.stderr)
    t0 = time.time()
    if args.impl == "ours":
        out = run_ours(args, rank, local, world)
        if not args.no_baseline:
            out.update(measure_baselines(args, out, rank, world))
    else:
        out = run_ddp(args, rank, local, world, args.ddp_mode)
    out["wall_s"] = time.time() - t0
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
