"""utils/numa.py: never raises, never widens the affinity, reports what it did (CPU-only box: NVML has nothing to say)."""
import os

from fpga_ai_nic_b200.utils import numa


def test_bind_is_harmless_without_nvml_affinity(monkeypatch):
    before = os.sched_getaffinity(0)
    info = numa.bind_to_gpu(0)
    assert set(info) >= {"bound", "cpus", "reason"}
    assert os.sched_getaffinity(0) <= before
    if not info["bound"]:
        assert os.sched_getaffinity(0) == before and info["reason"]


def test_bind_intersects_with_allowed_cpus(monkeypatch):
    allowed = sorted(os.sched_getaffinity(0))
    monkeypatch.setattr(numa, "gpu_cpu_set", lambda i: {allowed[0], 10 ** 6})   # one real CPU + one that does not exist
    try:
        info = numa.bind_to_gpu(0)
        assert os.sched_getaffinity(0) == {allowed[0]}
        assert info["cpus"] == 1 and (info["bound"] or len(allowed) == 1)
        monkeypatch.setattr(numa, "gpu_cpu_set", lambda i: {10 ** 6})
        info = numa.bind_to_gpu(0)
        assert not info["bound"] and "outside" in info["reason"]
        monkeypatch.setenv("AINIC_NUMA_BIND", "0")
        assert numa.bind_to_gpu(0)["reason"] == "AINIC_NUMA_BIND=0"
    finally:
        os.sched_setaffinity(0, allowed)
