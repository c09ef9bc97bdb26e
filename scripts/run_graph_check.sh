#!/bin/bash
# N-GPU check of graph replay with the device-side call number: process-per-GPU checks, then the
# reference config with and without graphs.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29516"
if [ -z "$SKIP_CHECK" ]; then
timeout 600 $TR scripts/mp_check.py > gpurun_out/mp_check_n$N.log 2>&1; echo "mp_check rc=$?"
grep -c '"ok": true' gpurun_out/mp_check_n$N.log; grep '"ok": false\|summary' gpurun_out/mp_check_n$N.log | cut -c1-300
fi
for g in ${GRAPHS:-1 0}; do
  AINIC_MLP_GRAPH=$g timeout 300 $TR bench.py --gpus $N --steps 200 --warmup 10 2> gpurun_out/bench_n${N}_g$g.err | grep '^{' > gpurun_out/bench_n${N}_g$g.json
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n${N}_g$g.json").readline())
    c=d["comm_counters"]; L=max(c["launches"],1); e=d["e2e"]
    print("n$N graph $g", round(d["ms_per_step"],4), round(d["value"]), "e2e", round(e["ms_per_step"],4), round(e["value"]), "bf16host", round(e["bf16_host_inputs"]["ms_per_step"],4), "percall_us", round(c["busy_ns"]/L/1e3,1), "b0", round(c["to_barrier0_ns"]/L/1e3,1), "b1", round(c["to_barrier1_ns"]/L/1e3,1), "timeouts", c["timeouts"], "loss", round(d["loss"],4), d["gpu_launches"], d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
except Exception as ex:
    print("n$N graph $g FAILED", ex)
PY
done
