"""The CLI-compatible CPU restatement of the reference application (BASELINE.json config 1:
"MLP data-parallel world_size=2 on CPU with host MPI allreduce").  Mode A (NIC all-reduce through
the fake device, overlapped with the backward pass) must train exactly like mode B (the reference's
commented-out MPI_Iallreduce + SGD path, sw/mlp_mpi_example_f32.cpp:590-680) when the wire format
is fp32, and the PERFDUMP line must have the reference's shape (:803-808)."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def app():
    from fpga_ai_nic_b200.utils import build
    outs = build.build_apps()
    assert outs
    return str(outs[0])


def _run(app, args, **env):
    e = dict(os.environ, OMP_NUM_THREADS="2", **env)
    p = subprocess.run([app] + [str(a) for a in args], capture_output=True, text=True, timeout=300, env=e)
    assert p.returncode == 0, p.stderr
    return p.stdout


def _field(out, name):
    m = re.search((r"^loss" if name == "loss" else name) + r"\s*=\s*([-0-9.e+]+)", out, re.M)
    assert m, out
    return float(m.group(1))


def test_mode_a_equals_mode_b_with_fp32_wire(app):
    args = [3, 64, 0, "A", 32, 32, 32, 128, 128, 128]
    a = _run(app, args, AINIC_CODEC="f32", AINIC_RANKS="2")
    args[3] = "B"
    b = _run(app, args, AINIC_RANKS="2")
    assert _field(a, "weight checksum") == _field(b, "weight checksum")
    assert _field(a, "loss") == _field(b, "loss")
    assert re.search(r"^PERFDUMP,BP,[^,]+,2,32,128,128,[0-9.]+,[0-9.]+$", a, re.M)
    for key in ("GFLOP ", "fp time", "GFLOPS ", "FC time compute/loss", "Bwdupd compute FIRST time overlaped"):
        assert key in a


@pytest.mark.parametrize("ranks,algo,codec", [(3, "ring", "bfp"), (2, "twoshot", "mxfp8"), (1, "auto", "bfp")])
def test_compressed_modes_train(app, ranks, algo, codec):
    out = _run(app, [4, 96, 3, "A", 64, 64, 64, 128, 256, 128], AINIC_RANKS=str(ranks), AINIC_ALGO=algo, AINIC_CODEC=codec,
               AINIC_LR="2.0")
    ref = _run(app, [4, 96, 3, "B", 64, 64, 64, 128, 256, 128], AINIC_RANKS=str(ranks), AINIC_LR="2.0")
    # compressed gradients follow the uncompressed run closely
    assert abs(_field(out, "loss") - _field(ref, "loss")) < 0.02 * abs(_field(ref, "loss"))
    assert abs(_field(out, "weight checksum") - _field(ref, "weight checksum")) < 0.05 * abs(_field(ref, "weight checksum")) + 0.05


def test_forward_only_and_usage(app):
    out = _run(app, [2, 32, 1, "F", 64, 64, 64, 64, 64])
    assert "PERFDUMP" in out
    assert "Usage" in _run(app, ["-h"])


def test_check_mode_reports_replica_state(app):
    """CHECK=1: two-shot keeps the replicas bit-identical; the reference-faithful ring lets the owner of a
    slice keep the unquantised sum while everybody else decodes BFP, so replicas differ."""
    args = [3, 96, 0, "A", 32, 32, 32, 128, 256, 128]
    out = _run(app, args, AINIC_RANKS="3", AINIC_ALGO="twoshot", CHECK="1")
    assert "CHECK: weight replicas bit-identical across 3 ranks" in out
    assert len(re.findall(r"^CHECK layer \d+: L1 norm of weights = [0-9.e+-]+, replicas identical$", out, re.M)) == 2
    out = _run(app, args, AINIC_RANKS="3", AINIC_ALGO="ring", CHECK="1")
    assert "CHECK: weight replicas differ across 3 ranks" in out
    out = _run(app, args, AINIC_RANKS="3", AINIC_ALGO="ring", AINIC_CODEC="f32", CHECK="1")
    assert "CHECK: weight replicas bit-identical across 3 ranks" in out
