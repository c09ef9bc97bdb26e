#!/bin/bash
# 8-GPU run: process-per-GPU checks, then the gradient-bucket sweep on the reference config.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518"
timeout 600 $TR scripts/mp_check.py > gpurun_out/mp_check_n$N.log 2>&1; echo "mp_check rc=$?"
grep -c '"ok": true' gpurun_out/mp_check_n$N.log; grep '"ok": false\|summary' gpurun_out/mp_check_n$N.log | cut -c1-300
for b in ${BUCKETS:-0 48 80 160}; do
  timeout 300 $TR bench.py --gpus $N --steps 200 --warmup 10 --no-e2e --bucket-mb $b 2> gpurun_out/bench_n${N}_b$b.err | grep '^{' > gpurun_out/bench_n${N}_b$b.json
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n${N}_b$b.json").readline())
c=d["comm_counters"]; L=max(c["launches"],1)
print("n$N bucket", "$b", round(d["ms_per_step"],4), round(d["value"]), "percall_us", round(c["busy_ns"]/L/1e3,1), "wait", round(c["wait_ns"]/L/1e3,1), "b0", round(c["to_barrier0_ns"]/L/1e3,1), "b1", round(c["to_barrier1_ns"]/L/1e3,1), "launches", L, d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
PY
done
