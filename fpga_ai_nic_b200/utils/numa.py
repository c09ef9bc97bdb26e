"""Keep a rank's host side next to its GPU.

With one process per GPU, every rank copies its minibatch from pinned host memory each step.  Pinned
pages are placed on the NUMA node of the thread that allocates them; when all ranks of a box sit on one
socket, eight simultaneous host -> device copies share that socket's memory controllers and, for half of
the GPUs, the inter-socket link (measured on the 8-GPU boxes of this pool: 24 GB/s per GPU with eight
copies in flight against 55 GB/s alone - which bound the end-to-end step).  ``bind_to_gpu`` restricts the
calling process to the CPUs NVML reports as local to the GPU, BEFORE any pinned allocation is made, so
the default first-touch policy puts the staging buffers on the right node.

The reference pins its MPI ranks the same way from the launcher (``sw/run.sh``: ``I_MPI_PIN_DOMAIN`` /
one rank per FPGA host); this is the one-box equivalent.
"""
from __future__ import annotations

import os

from .clocks import _physical_index


def gpu_cpu_set(local_index: int) -> set[int]:
    """CPUs local to the GPU (NVML ideal affinity), empty when NVML cannot tell."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(_physical_index(local_index))
        ncpu = os.cpu_count() or 1
        words = (ncpu + 63) // 64
        mask = nv.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        return {c for c in cpus if c < ncpu}
    except Exception:
        return set()


def bind_to_gpu(local_index: int) -> dict:
    """Restrict this process to the CPUs next to GPU ``local_index`` (intersected with what the container
    allows).  Returns what was done; never raises.  ``AINIC_NUMA_BIND=0`` switches it off."""
    info = {"bound": False, "cpus": 0, "reason": ""}
    if os.environ.get("AINIC_NUMA_BIND", "1") == "0":
        info["reason"] = "AINIC_NUMA_BIND=0"
        return info
    if not hasattr(os, "sched_setaffinity"):
        info["reason"] = "no sched_setaffinity"
        return info
    want = gpu_cpu_set(local_index)
    if not want:
        info["reason"] = "NVML reports no CPU affinity"
        return info
    try:
        allowed = os.sched_getaffinity(0)
        use = want & allowed
        if not use:
            info["reason"] = "GPU-local CPUs are outside this process's cpuset"
            return info
        if use == allowed:
            info["reason"] = "already confined to GPU-local CPUs"
            info["cpus"] = len(use)
            return info
        os.sched_setaffinity(0, use)
        info.update(bound=True, cpus=len(use))
    except OSError as ex:
        info["reason"] = f"sched_setaffinity: {ex}"
    return info
