#!/bin/bash
# 2-GPU validation of gradient bucketing: unit tests, process-per-GPU checks, bucket sweep.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_trainer.py -m gpu -x -q > gpurun_out/pytest_trainer.log 2>&1; echo "pytest trainer rc=$?"
tail -3 gpurun_out/pytest_trainer.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
timeout 600 $TR scripts/mp_check.py > gpurun_out/mp_check_n2.log 2>&1; echo "mp_check rc=$?"
grep -c '"ok": true' gpurun_out/mp_check_n2.log; grep '"ok": false\|summary\|bucketed' gpurun_out/mp_check_n2.log | cut -c1-300
for b in 0 32 80 1000; do
  timeout 300 $TR bench.py --gpus 2 --steps 200 --warmup 10 --no-e2e --bucket-mb $b 2> gpurun_out/bench_n2_b$b.err | grep '^{' > gpurun_out/bench_n2_b$b.json
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n2_b$b.json").readline())
c=d["comm_counters"]; L=max(c["launches"],1)
print("n2 bucket", "$b", d["ms_per_step"], d["value"], "percall_us", c["busy_ns"]/L/1e3, "wait", c["wait_ns"]/L/1e3, "b0", c["to_barrier0_ns"]/L/1e3, "b1", c["to_barrier1_ns"]/L/1e3, "launches", L)
PY
done
