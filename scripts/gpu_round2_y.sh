#!/bin/bash
# Round-2 call y (N GPUs): end-to-end step with / without NUMA binding of the ranks (no DDP arms).
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
export AINIC_TIMEOUT_MS=5000
N=${1:-8}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
python - <<'PY'
import os
print("cpus", os.cpu_count(), "allowed", len(os.sched_getaffinity(0)))
try:
    import pynvml as nv
    nv.nvmlInit()
    for i in range(nv.nvmlDeviceGetCount()):
        h = nv.nvmlDeviceGetHandleByIndex(i)
        m = nv.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        print("gpu", i, [hex(int(x)) for x in m])
except Exception as e:
    print("nvml", e)
os.system("cat /sys/devices/system/node/online; nproc")
PY
timeout 300 $RUN --master-port 29721 bench.py --gpus $N --steps 20 --warmup 5 --no-baseline > gpurun_out/y_bench_bind_n$N.json 2> gpurun_out/y_bench_bind_n$N.err; echo "bench rc=$?"
AINIC_NUMA_BIND=0 timeout 300 $RUN --master-port 29722 bench.py --gpus $N --steps 20 --warmup 5 --no-baseline > gpurun_out/y_bench_nobind_n$N.json 2> gpurun_out/y_bench_nobind_n$N.err; echo "bench rc=$?"
python - <<PY
import json
for f in ('bind','nobind'):
    try:
        d=json.loads(open('gpurun_out/y_bench_%s_n$N.json' % f).read().strip().splitlines()[-1])
        e=d['e2e']
        print(f, 'dev', d.get('ms_per_step'), 'e2e', e.get('ms_per_step'), 'copy alone GB/s', e.get('h2d_copy_alone_gbps'), 'bf16 host', e.get('bf16_host_inputs',{}).get('ms_per_step'), e.get('numa_bind'))
    except Exception as ex: print('ERR', ex)
PY
