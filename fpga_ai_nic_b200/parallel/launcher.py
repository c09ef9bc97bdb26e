"""Launcher and sweep driver: the counterpart of sw/run.sh (mpirun -n 3 -ppn 1 ... plus a commented
6-node sweep over per-node minibatches with five repetitions logged to ./log/<nodes>_<mode>.log,
sw/run.sh:12-36) and of the ring-route set-up script sw/setup_route.sh (nothing to route here:
NVSwitch connects every GPU to every peer).

    python -m fpga_ai_nic_b200.parallel.launcher --gpus 8 -- bench.py --steps 50
    python -m fpga_ai_nic_b200.parallel.launcher --sweep --gpus-list 1,2,4,8 --mb-list 1792,448 --reps 5
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(gpus: int, script_args: list[str], env: dict | None = None, timeout: float | None = None) -> subprocess.CompletedProcess:
    """One process per GPU on this node; rendezvous on 127.0.0.1 (the container hostname may not resolve)."""
    e = dict(os.environ)
    e.update(env or {})
    e.setdefault("OMP_NUM_THREADS", "4")
    if gpus <= 1:
        cmd = [sys.executable] + script_args
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={gpus}", "--master-addr",
               "127.0.0.1", "--master-port", str(free_port())] + script_args
    return subprocess.run(cmd, capture_output=True, text=True, env=e, timeout=timeout)


def sweep(gpus_list, mb_list, layers: int, width: int, reps: int, impls, logdir: Path, steps: int, warmup: int) -> list:
    """sw/run.sh:17-36: for each node count and per-node minibatch, repeat and log."""
    logdir.mkdir(parents=True, exist_ok=True)
    results = []
    for impl in impls:
        for n in gpus_list:
            for mb in mb_list:
                log = logdir / f"{n}_{impl}_{mb}.log"
                with open(log, "a") as fh:
                    for rep in range(reps):
                        p = launch(n, [str(ROOT / "bench.py"), "--gpus", str(n), "--impl", impl, "--mb", str(mb), "--layers",
                                       str(layers), "--width", str(width), "--steps", str(steps), "--warmup", str(warmup),
                                       "--no-e2e"], timeout=1200)
                        fh.write(p.stdout + p.stderr)
                        line = next((l for l in p.stdout.splitlines() if l.startswith("{")), None)
                        if line:
                            r = json.loads(line)
                            results.append({"impl": impl, "gpus": n, "mb": mb, "rep": rep, "samples_per_s": r["value"],
                                            "ms_per_step": r["ms_per_step"]})
                            print(f"{impl:5s} gpus={n} mb={mb} rep={rep}: {r['value']:.0f} samples/s ({r['ms_per_step']:.3f} ms/step)",
                                  flush=True)
                        else:
                            print(f"{impl} gpus={n} mb={mb} rep={rep}: FAILED rc={p.returncode}", flush=True)
    return results


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--gpus-list", default="1,2,4,8")
    ap.add_argument("--mb-list", default="1792,448")
    ap.add_argument("--layers", type=int, default=20, help="sw/run.sh sweep uses 20 layers")
    ap.add_argument("--width", type=int, default=2048)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--impls", default="ours,ddp")
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--logdir", default=str(ROOT / "log"))
    ap.add_argument("rest", nargs=argparse.REMAINDER)
    a = ap.parse_args()
    if a.sweep:
        res = sweep([int(x) for x in a.gpus_list.split(",")], [int(x) for x in a.mb_list.split(",")], a.layers, a.width,
                    a.reps, a.impls.split(","), Path(a.logdir), a.steps, a.warmup)
        print(json.dumps(res))
        return 0
    rest = a.rest[1:] if a.rest and a.rest[0] == "--" else a.rest
    if not rest:
        ap.error("nothing to launch")
    p = launch(a.gpus, rest)
    sys.stdout.write(p.stdout)
    sys.stderr.write(p.stderr)
    return p.returncode


if __name__ == "__main__":
    sys.exit(main())
