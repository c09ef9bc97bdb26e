"""CPU checks of the host-side utilities: the reference's report format (PERFDUMP line and FLOP
accounting, sw/mlp_mpi_example_f32.cpp:794-808), the launcher plumbing and the MLP config math."""
import re
import subprocess
import sys

from fpga_ai_nic_b200.models.mlp import MLPConfig
from fpga_ai_nic_b200.parallel import launcher
from fpga_ai_nic_b200.utils import profiling


def test_flop_accounting_matches_reference_formula():
    sizes = [2048] * 11
    g = profiling.flops_reference(sizes, 5376, 1)           # 3 ranks x 1792 (sw/run.sh:16)
    want = (9 * 6.0 + 4.0) * 5376 * 2048 * 2048 / 1e9
    assert abs(g - want) < 1e-6 * want
    cfg = MLPConfig.reference()
    assert abs(cfg.flops_per_step(5376) / 1e9 - want) < 1e-6 * want
    assert cfg.param_count() == 10 * 2048 * 2048 and cfg.nlayers == 10 and cfg.mb == 1792


def test_perfdump_block_has_reference_shape():
    txt = profiling.perfdump([256, 512, 128], 64, 2, 10, 0.5, {"fwd_fc": 100.0, "bwdupd_fc": 200.0, "bwdupd_first": 50.0})
    lines = txt.splitlines()
    assert lines[0].startswith("GFLOP  = ") and lines[1].startswith("fp time = ") and lines[2].startswith("GFLOPS  = ")
    assert re.match(r"^PERFDUMP,BP,[^,]+,2,64,256,512,[0-9.]+,[0-9.]+$", lines[3])
    assert any(l.startswith("Bwdupd compute FIRST time overlaped") for l in lines)
    assert "samples/s = 2560" in txt


def test_launcher_runs_a_script_single_process(tmp_path):
    script = tmp_path / "hello.py"
    script.write_text("import os; print('rank', os.environ.get('RANK', '0'))\n")
    p = launcher.launch(1, [str(script)])
    assert p.returncode == 0 and "rank 0" in p.stdout


def test_launcher_two_ranks_gloo(tmp_path):
    script = tmp_path / "two.py"
    script.write_text(
        "import os, sys, torch, torch.distributed as dist\n"
        "dist.init_process_group('gloo')\n"
        "t = torch.tensor([float(dist.get_rank() + 1)])\n"
        "dist.all_reduce(t)\n"
        "open(os.path.join(sys.argv[1], 'rank%d.txt' % dist.get_rank()), 'w').write(str(t.item()))\n"
        "dist.destroy_process_group()\n")
    p = launcher.launch(2, [str(script), str(tmp_path)], timeout=120)
    assert p.returncode == 0, p.stderr[-2000:]
    # results go through files: torchrun does not always forward the last lines of a worker's stdout
    assert [(tmp_path / f"rank{r}.txt").read_text() for r in range(2)] == ["3.0", "3.0"]


def test_bench_reference_arm_reports_unavailable():
    p = subprocess.run([sys.executable, "bench.py", "--impl", "reference"], capture_output=True, text=True, cwd=launcher.ROOT)
    assert p.returncode == 0 and '"impl": "reference"' in p.stdout and "unavailable" in p.stdout


def test_bucket_plan_and_heap_sizing(monkeypatch):
    """Gradient buckets: per layer on one rank (the reference's behaviour, :754), 80 MiB with peers,
    explicit / environment overrides; the heap must hold the staging of the largest bucket."""
    monkeypatch.delenv("AINIC_BUCKET_MB", raising=False)
    cfg = MLPConfig.reference()                                    # 10 layers x 16 MiB
    assert cfg.bucket_plan(1) == [(l, l) for l in range(9, -1, -1)]
    assert cfg.bucket_plan(8) == [(5, 9), (0, 4)]
    assert MLPConfig(sizes=[2048] * 11, bucket_mb=48).bucket_plan(8) == [(7, 9), (4, 6), (1, 3), (0, 0)]
    assert MLPConfig(sizes=[2048] * 11, bucket_mb=1e6).bucket_plan(2) == [(0, 9)]
    assert MLPConfig(sizes=[8192] * 9).bucket_plan(8) == [(l, l) for l in range(7, -1, -1)]   # 256 MiB layers
    monkeypatch.setenv("AINIC_BUCKET_MB", "32")
    assert cfg.bucket_plan(1) == [(8, 9), (6, 7), (4, 5), (2, 3), (0, 1)]
    assert MLPConfig(sizes=[2048] * 11, bucket_mb=0).bucket_plan(8) == [(l, l) for l in range(9, -1, -1)]
    monkeypatch.delenv("AINIC_BUCKET_MB")
    # every plan covers each layer exactly once, top down
    for world, mb in ((1, None), (2, None), (8, 20.0), (8, 100.0)):
        plan = MLPConfig(sizes=[256, 512, 384, 128, 640], bucket_mb=mb).bucket_plan(world)
        covered = [l for lo, hi in plan for l in range(hi, lo - 1, -1)]
        assert covered == list(range(3, -1, -1))
    p = cfg.param_count()
    assert cfg.heap_bytes(1) >= p * 8 and cfg.heap_bytes(8) > cfg.heap_bytes(1)
    assert MLPConfig(sizes=[2048] * 11, bucket_mb=1e6).heap_bytes(8) > cfg.heap_bytes(8)


def test_native_bucket_planner_matches_the_python_mirror(monkeypatch):
    """csrc/api/bucket_plan.h is what the GPU trainer runs; the CPU library exports it, so the plan the
    trainer will use is checked here against MLPConfig.bucket_plan for thresholds and explicit schedules."""
    import random
    from fpga_ai_nic_b200.ops import oracle as O
    rng = random.Random(7)
    monkeypatch.delenv("AINIC_BUCKET_MB", raising=False)
    for trial in range(200):
        nl = rng.randint(1, 12)
        sizes = [rng.choice([64, 128, 256, 512, 1024, 2048]) for _ in range(nl + 1)]
        layer_bytes = [sizes[l] * sizes[l + 1] * 4 for l in range(nl)]
        if trial % 2:
            sched = [rng.randint(1, 5) for _ in range(rng.randint(1, 4))]
            monkeypatch.setenv("AINIC_BUCKET_LAYERS", ",".join(map(str, sched)))
            want = MLPConfig(sizes=sizes).bucket_plan(8)
            got = O.bucket_plan(layer_bytes, 0, sched)
        else:
            monkeypatch.delenv("AINIC_BUCKET_LAYERS", raising=False)
            mb = rng.choice([0.0, 0.01, 0.5, 3.0, 20.0, 1e6])
            want = MLPConfig(sizes=sizes, bucket_mb=mb).bucket_plan(8)
            got = O.bucket_plan(layer_bytes, int(mb * (1 << 20)), None)
        assert got == want, (sizes, got, want)
        covered = [l for lo, hi in got for l in range(hi, lo - 1, -1)]
        assert covered == list(range(nl - 1, -1, -1))
    monkeypatch.setenv("AINIC_BUCKET_LAYERS", "4,3,2,1")
    assert MLPConfig.reference().bucket_plan(8) == [(6, 9), (3, 5), (1, 2), (0, 0)]
    assert O.bucket_plan([16 << 20] * 10, 0, [4, 3, 2, 1]) == [(6, 9), (3, 5), (1, 2), (0, 0)]
