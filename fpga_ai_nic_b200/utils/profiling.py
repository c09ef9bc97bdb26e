"""Per-phase profiling and the reference's report format.

The reference accumulates thread-0 timer deltas for forward FC, forward loss, backward loss and
backward+update FC, plus the first (non-overlapped) backward layer, and prints GFLOP / fp time /
GFLOPS / a machine-readable PERFDUMP line (sw/mlp_mpi_example_f32.cpp:699-751, :794-816).
``PhaseTimer`` does the same with CUDA events on the launching stream (never wall clock);
``perfdump`` prints the same block.
"""
from __future__ import annotations

from collections import defaultdict
from contextlib import contextmanager

import torch


class PhaseTimer:
    def __init__(self, enabled: bool = True):
        self.enabled = enabled
        self._pending = defaultdict(list)
        self.totals_ms = defaultdict(float)
        self.counts = defaultdict(int)

    @contextmanager
    def phase(self, name: str, stream=None):
        if not self.enabled:
            yield
            return
        s = stream or torch.cuda.current_stream()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s)
        try:
            yield
        finally:
            b.record(s)
            self._pending[name].append((a, b))

    def collect(self) -> dict:
        torch.cuda.synchronize()
        for name, evs in self._pending.items():
            for a, b in evs:
                self.totals_ms[name] += a.elapsed_time(b)
                self.counts[name] += 1
        self._pending.clear()
        return dict(self.totals_ms)


def flops_reference(sizes, global_mb: int, iters: int) -> float:
    """6*MB*C*K per layer, 4* for layer 0 (sw/mlp_mpi_example_f32.cpp:794-798), in GFLOP."""
    g = 0.0
    L = len(sizes) - 1
    for l in range(L - 1, 0, -1):
        g += 6.0 * global_mb * sizes[l] * sizes[l + 1] * iters / 1e9
    g += 4.0 * global_mb * sizes[0] * sizes[1] * iters / 1e9
    return g


def perfdump(sizes, mb_per_rank: int, nranks: int, iters: int, seconds: float, phases_ms: dict | None = None,
             version: str = "b200-ainic") -> str:
    gflop = flops_reference(sizes, mb_per_rank * nranks, iters)
    lines = [f"GFLOP  = {gflop / iters:.5g}", f"fp time = {seconds / iters:.5g}", f"GFLOPS  = {gflop / seconds:.5g}",
             "PERFDUMP,BP,%s,%i,%i,%s,%f,%f" % (version, nranks, mb_per_rank, ",".join(str(c) for c in sizes[:-1]),
                                               seconds / iters, gflop / seconds)]
    if phases_ms:
        tot = sum(phases_ms.get(k, 0.0) for k in ("fwd_fc", "fwd_loss", "bwd_loss", "bwdupd_fc")) / 1e3
        lines.append(f"FC time compute/loss = {tot / iters:.5g}")
        first = phases_ms.get("bwdupd_first", 0.0) / 1e3
        lines.append(f"Bwdupd compute FIRST time overlaped = {first / iters:.5g}")
        lines.append(f"Bwdupd compute time overlaped = {(phases_ms.get('bwdupd_fc', 0.0) / 1e3 - first) / iters:.5g}")
        if "allreduce" in phases_ms:
            lines.append(f"Allreduce+update (comm stream) = {phases_ms['allreduce'] / 1e3 / iters:.5g}")
    lines.append(f"samples/s = {mb_per_rank * nranks * iters / seconds:.6g}")
    return "\n".join(lines)
