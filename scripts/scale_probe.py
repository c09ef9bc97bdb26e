#!/usr/bin/env python
"""Several trainer configurations in ONE process group (one rendezvous, one import per rank): every
configuration is a set of AINIC_* environment overrides read when the trainer is created.  Prints one
JSON line per configuration on rank 0: device-timed ms/step (max over ranks), per-step times, the
all-reduce kernels' own phase counters.  Used to compare flow / bucket schedules / fused epilogues on
multi-GPU boxes without paying the start-up cost per configuration.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/scale_probe.py --configs 'AINIC_FLOW=1' 'AINIC_FLOW=0' 'AINIC_FLOW=0,AINIC_BUCKET_LAYERS=4:3:2:1'
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", nargs="+", default=["AINIC_FLOW=1", "AINIC_FLOW=0"])
    ap.add_argument("--layers", type=int, default=10)
    ap.add_argument("--width", type=int, default=2048)
    ap.add_argument("--mb", type=int, default=1792)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--repeats", type=int, default=3)
    ap.add_argument("--codec", default="auto")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from fpga_ai_nic_b200.models.mlp import MLPConfig, MLPTrainer
    from fpga_ai_nic_b200.ops import kernels as K
    g = torch.Generator().manual_seed(1234 + rank)
    xs = [K.cast_bf16((torch.rand(args.mb, args.width, generator=g) * 0.1 - 0.05).cuda()) for _ in range(4)]
    ys = [torch.randint(0, args.width, (args.mb,), generator=g, dtype=torch.int32).cuda() for _ in range(4)]
    results = []
    for spec in args.configs:
        over = dict(kv.split("=", 1) for kv in spec.split(",") if kv)
        over = {k: v.replace(":", ",") for k, v in over.items()}
        saved = {k: os.environ.get(k) for k in over}
        os.environ.update(over)
        try:
            cfg = MLPConfig.reference(args.layers, args.width, args.mb)
            cfg.codec = over.get("CODEC", args.codec)
            tr = MLPTrainer(cfg)
            reps = []
            for rep in range(args.repeats):
                for i in range(args.warmup if rep == 0 else 2):
                    tr.step(xs[i % 4], ys[i % 4])
                if world > 1:
                    dist.barrier()
                torch.cuda.synchronize()
                c0 = tr.comm.counters()
                ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
                ev[0].record()
                for i in range(args.steps):
                    tr.step(xs[i % 4], ys[i % 4])
                    ev[i + 1].record()
                tr.wait()
                torch.cuda.synchronize()
                c1 = tr.comm.counters()
                t = torch.tensor([ev[0].elapsed_time(ev[-1])], device="cuda", dtype=torch.float64)
                if world > 1:
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                n = max(1, c1["launches"] - c0["launches"])
                reps.append({"ms_per_step": float(t.item()) / args.steps,
                             "step_ms": [round(ev[i].elapsed_time(ev[i + 1]), 3) for i in range(args.steps)],
                             "calls": n, "busy_us": (c1["busy_ns"] - c0["busy_ns"]) / n / 1e3,
                             "wait_us": (c1["wait_ns"] - c0["wait_ns"]) / n / 1e3,
                             "b0_us": (c1["to_barrier0_ns"] - c0["to_barrier0_ns"]) / n / 1e3,
                             "b1_us": (c1["to_barrier1_ns"] - c0["to_barrier1_ns"]) / n / 1e3,
                             "timeouts": c1["timeouts"]})
            # replicas identical?
            sig = torch.stack([tr.buffer("weight", l).view(torch.int32).to(torch.int64).sum() for l in range(cfg.nlayers)])
            same = True
            if world > 1:
                allsig = [torch.zeros_like(sig) for _ in range(world)]
                dist.all_gather(allsig, sig)
                same = all(torch.equal(allsig[0], s) for s in allsig)
            loss = float(tr.loss_dev.item()) / cfg.mb
            trace = tr.comm.flow_trace(cfg.nlayers) if os.environ.get("AINIC_FLOW_TRACE") else None
            res = {"config": spec, "world": world, "best_ms": min(r["ms_per_step"] for r in reps), "reps": reps,
                   "replicas_identical": same, "loss": loss, "codec": tr.codec, "flow_trace": trace}
            tr.close()
            tr.comm.close()
        except Exception as ex:  # keep probing
            res = {"config": spec, "world": world, "error": f"{type(ex).__name__}: {str(ex)[:300]}"}
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
        torch.cuda.empty_cache()
        results.append(res)
        if rank == 0:
            brief = dict(res)
            if "reps" in brief:
                brief["reps"] = [{k: (v if k != "step_ms" else None) for k, v in r.items()} for r in brief["reps"]]
            print(json.dumps(brief), flush=True)
    if rank == 0 and args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            json.dump(results, f, indent=1)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
