#!/bin/bash
# N-GPU validation of the GEMM -> reduce-scatter epilogue: kernel test, process-per-GPU checks, mode sweep.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
if [ -z "$SKIP_PYTEST" ]; then
python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "scatter or compressed" > gpurun_out/pytest_scatter.log 2>&1; echo "pytest scatter rc=$?"
tail -3 gpurun_out/pytest_scatter.log
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514"
if [ -z "$SKIP_CHECK" ]; then
timeout 600 $TR scripts/mp_check.py > gpurun_out/mp_check_n$N.log 2>&1; echo "mp_check rc=$?"
grep -c '"ok": true' gpurun_out/mp_check_n$N.log; grep '"ok": false\|summary' gpurun_out/mp_check_n$N.log | cut -c1-300
fi
for cfg in ${CFGS:-0:0 0:80 1:80 2:80 2:0}; do
  f=${cfg%%:*}; b=${cfg##*:}
  AINIC_FUSE_COMPRESS=$f timeout 300 $TR bench.py --gpus $N --steps 200 --warmup 10 --no-e2e --bucket-mb $b 2> gpurun_out/bench_n${N}_f${f}_b$b.err | grep '^{' > gpurun_out/bench_n${N}_f${f}_b$b.json
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n${N}_f${f}_b$b.json").readline())
    c=d["comm_counters"]; L=max(c["launches"],1)
    print("n$N fuse $f bucket $b", round(d["ms_per_step"],4), round(d["value"]), "percall_us", round(c["busy_ns"]/L/1e3,1), "wait", round(c["wait_ns"]/L/1e3,1), "b0", round(c["to_barrier0_ns"]/L/1e3,1), "b1", round(c["to_barrier1_ns"]/L/1e3,1), "launches", L, "loss", round(d["loss"],4), d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
except Exception as e:
    print("n$N fuse $f bucket $b FAILED", e)
PY
done
