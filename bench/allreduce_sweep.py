"""Fused all-reduce + weight-update bandwidth sweep (BASELINE.json config 3): 1 KB - 1 GB at N GPUs,
{f32, bf16, reference-BFP, MXFP8} x {oneshot, twoshot, ring} vs ncclAllReduce + a separate SGD kernel.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29520 bench/allreduce_sweep.py --max-mb 1024 --out gpurun_out/sweep_n8.json

Timing: CUDA events around every single call on the launching stream, L2 flushed (256 MiB memset)
between calls, >= 3 warm-ups, median over iterations, max over ranks.  Reported per row:
  algbw  = S / t                      (S = fp32 bytes of the bucket)
  busbw  = S * 2(N-1)/N / t           (NCCL's bus-bandwidth convention)
  wire   = bytes each GPU actually moves over NVLink per direction (compressed)
  roofline fraction = max(wire / 770 GB/s, HBM bytes / measured HBM peak) / t
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fpga_ai_nic_b200.ops import kernels as K  # noqa: E402
from fpga_ai_nic_b200.parallel.comm import Communicator  # noqa: E402

NVLINK_GBS = 770.0     # measured peer copy per direction (B200_PROFILING.md)
RATIO = {"f32": 1.0, "bf16": 0.5, "bfp": 544 / 2048, "mxfp8": 528 / 2048}


def hbm_peak():
    try:
        with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-kb", type=int, default=1)
    ap.add_argument("--max-mb", type=int, default=256)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    ap.add_argument("--codecs", default="f32,bf16,bfp,mxfp8")
    ap.add_argument("--algos", default="oneshot,twoshot,ring")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hbm, hbm_src = hbm_peak()
    max_count = (args.max_mb << 20) // 4
    comm = Communicator(max_count * 4 * 2 + (max_count // 512 + 1) * (2 * 1024 + 2 * 2048) + (64 << 20))
    g = comm.empty(max_count)
    w = comm.empty(max_count)
    sh = torch.empty(max_count, dtype=torch.bfloat16, device="cuda")
    g.normal_(0, 1e-3)
    w.normal_(0, 1)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    # staging inside the communicator grows but is never returned: size it once for the largest bucket
    comm.allreduce_fused(g, w, 0.0, sh, codec=K.CODEC_BF16, algo=K.ALGO_TWOSHOT)
    torch.cuda.synchronize()

    def timed(fn, iters):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(iters):
            flush.zero_()
            if world > 1:
                dist.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        ts.sort()
        t = torch.tensor([ts[len(ts) // 2]], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    g_nvls = None
    if world > 1 and comm.enable_nvls(max_count * 6 + (8 << 20)):
        g_nvls = comm.nvls_empty(max_count)
        g_nvls.normal_(0, 1e-3)
        g_nvls16 = comm.nvls_empty(max_count, torch.bfloat16)
        g_nvls16.normal_(0, 1e-3)
    elif rank == 0:
        print("# NVLS unavailable:", getattr(comm, "nvls_error", ""))

    size = args.min_kb << 10
    while size <= (args.max_mb << 20):
        count = size // 4
        gv, wv, sv = g[:count], w[:count], sh[:count]
        iters = args.iters if size <= (64 << 20) else max(3, args.iters // 2)

        def nccl():
            if world > 1:
                dist.all_reduce(gv)
            wv.add_(gv, alpha=-0.1)

        ms = timed(nccl, iters)
        rows.append({"impl": "nccl+sgd", "bytes": size, "ms": ms})
        for codec in args.codecs.split(","):
            for algo in args.algos.split(","):
                if algo == "ring" and codec not in ("f32", "bfp"):
                    continue
                if algo == "oneshot" and size > (8 << 20) and world > 2:
                    continue
                if algo == "ring" and size > (256 << 20):
                    continue
                fn = lambda: comm.allreduce_fused(gv, wv, 0.1, sv, codec=K.CODEC_NAMES[codec], bfp_mode=K.CLEAN,
                                                  algo=K.ALGO_NAMES[algo])
                try:
                    ms = timed(fn, iters)
                except Exception as ex:  # keep sweeping
                    rows.append({"impl": f"{algo}/{codec}", "bytes": size, "error": str(ex)})
                    continue
                r = RATIO[codec]
                wire = size * r * ((world - 1) if algo == "oneshot" else 2 * (world - 1) / world)
                hbm_bytes = size * (3 + 0.5 + (r if codec != "f32" else 0) * 2) + (size if codec != "f32" else 0)
                roof_ms = max(wire / (NVLINK_GBS * 1e6), hbm_bytes / (hbm * 1e6))
                rows.append({"impl": f"{algo}/{codec}", "bytes": size, "ms": ms, "wire_bytes": wire,
                             "roofline_ms": roof_ms, "roofline_frac": roof_ms / ms})
        if g_nvls is not None and count % 512 == 0:
            gn = g_nvls[:count]
            ms = timed(lambda: comm.allreduce_nvls(gn, wv, 0.1, sv), iters)
            wire = size * 1.0          # ~S per direction: ld_reduce pulls (N-1)/N*S, multimem.st pushes S/N
            hbm_bytes = size * 3.5
            roof_ms = max(wire / (NVLINK_GBS * 1e6), hbm_bytes / (hbm * 1e6))
            rows.append({"impl": "nvls/f32", "bytes": size, "ms": ms, "wire_bytes": wire, "roofline_ms": roof_ms,
                         "roofline_frac": roof_ms / ms})
            # bf16 bucket.  "nvls/bf16" = the collective alone (the producer already wrote bf16, as a GEMM epilogue
            # would); "cast+nvls/bf16" = fp32 gradient -> bf16 cast kernel -> collective, comparable with the fused
            # codecs above, which also start from the fp32 gradient.
            gb = g_nvls16[:count]
            ms = timed(lambda: comm.allreduce_nvls(gb, wv, 0.1, sv), iters)
            wire = size * 0.5
            hbm_bytes = size * (0.5 * 2 + 2 + 0.5)      # bf16 bucket read + multicast write, fp32 weights r/w, bf16 shadow
            roof_ms = max(wire / (NVLINK_GBS * 1e6), hbm_bytes / (hbm * 1e6))
            rows.append({"impl": "nvls/bf16", "bytes": size, "ms": ms, "wire_bytes": wire, "roofline_ms": roof_ms,
                         "roofline_frac": roof_ms / ms})

            def cast_nvls():
                K.cast_bf16(gv, out=gb)
                comm.allreduce_nvls(gb, wv, 0.1, sv)

            ms = timed(cast_nvls, iters)
            hbm_bytes = size * (1 + 0.5 + 0.5 * 2 + 2 + 0.5)
            roof_ms = max(wire / (NVLINK_GBS * 1e6), hbm_bytes / (hbm * 1e6))
            rows.append({"impl": "cast+nvls/bf16", "bytes": size, "ms": ms, "wire_bytes": wire, "roofline_ms": roof_ms,
                         "roofline_frac": roof_ms / ms})
        size *= 4
    for r in rows:
        if "ms" in r:
            r["algbw_gbs"] = r["bytes"] / r["ms"] / 1e6
            r["busbw_gbs"] = r["algbw_gbs"] * (2 * (world - 1) / world if world > 1 else 1.0)
    ctr = comm.counters()
    if rank == 0:
        out = {"world": world, "hbm_peak_gbs": hbm, "hbm_peak_source": hbm_src, "nvlink_gbs_per_dir": NVLINK_GBS,
               "timeouts": ctr["timeouts"], "rows": rows}
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            with open(args.out, "w") as f:
                json.dump(out, f, indent=1)
        print(f"# world={world}  (ms; busbw GB/s; fraction of roofline)  timeouts={ctr['timeouts']}")
        sizes = sorted(set(r["bytes"] for r in rows))
        impls = []
        for r in rows:
            if r["impl"] not in impls:
                impls.append(r["impl"])
        print("bytes".rjust(12) + "".join(i.rjust(22) for i in impls))
        for s in sizes:
            line = str(s).rjust(12)
            for i in impls:
                m = [r for r in rows if r["bytes"] == s and r["impl"] == i and "ms" in r]
                if m:
                    f = m[0].get("roofline_frac")
                    line += (f"{m[0]['ms']:.3f}/{m[0]['busbw_gbs']:.0f}" + (f"/{f:.2f}" if f else "")).rjust(22)
                else:
                    line += "-".rjust(22)
            print(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
