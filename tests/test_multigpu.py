"""Process-per-GPU checks (scripts/mp_check.py): world 1 always, world 2/4/8 when the box has the GPUs.
The multi-rank runs exercise the real path: CUDA IPC symmetric heap, peer loads/stores over NVLink,
system-scope flags - against the CPU oracle, bit for bit."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(world: int):
    env = dict(os.environ, AINIC_TIMEOUT_MS="20000")
    if world == 1:
        cmd = [sys.executable, os.path.join(ROOT, "scripts", "mp_check.py")]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", str(29600 + world), os.path.join(ROOT, "scripts", "mp_check.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    lines = [json.loads(l) for l in p.stdout.splitlines() if l.startswith("{")]
    assert p.returncode == 0, (p.stdout[-3000:], p.stderr[-3000:])
    summary = [l for l in lines if l.get("check") == "summary"]
    assert summary and summary[0]["failures"] == [], summary
    assert all(l.get("ok", True) for l in lines)
    return lines


def test_process_per_gpu_world1():
    _run(1)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_process_per_gpu_multi(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    _run(world)
