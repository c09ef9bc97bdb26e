#!/bin/bash
# Round-2 call ag (N GPUs): deep / wide models through bench.py, i.e. with the PyTorch DDP arms in the same invocation.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_NO_BUILD=1
export AINIC_TIMEOUT_MS=5000
N=${1:-1}; WHAT=${2:-deep,wide}
if [ "$N" = "1" ]; then RUN="python"; else RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29731"; fi
if [[ $WHAT == *deep* ]]; then
timeout 900 $RUN bench.py --gpus $N --steps 20 --warmup 5 --layers 8 --width 8192 > gpurun_out/ag_deep_n$N.json 2> gpurun_out/ag_deep_n$N.err; echo "deep rc=$?"
fi
if [[ $WHAT == *wide* ]]; then
timeout 1200 $RUN bench.py --gpus $N --steps 10 --warmup 3 --layers 16 --width 16384 > gpurun_out/ag_wide_n$N.json 2> gpurun_out/ag_wide_n$N.err; echo "wide rc=$?"
fi
python - <<PY
import json
for c in ('deep','wide'):
    try:
        d=json.loads(open('gpurun_out/ag_%s_n$N.json' % c).read().strip().splitlines()[-1])
        b=d.get('baseline',{}).get('ddp',{})
        print(c, 'ours', d.get('ms_per_step'), 'e2e', d['e2e'].get('ms_per_step'), 'vs', d.get('vs_baseline'), {k:(v.get('ms_per_step'), v.get('cuda_graph')) for k,v in b.items() if isinstance(v,dict)}, d.get('correctness'))
    except Exception as e: print(c, 'ERR', e)
PY
for f in gpurun_out/ag_*_n$N.err; do tail -n 3 $f | cut -c1-300; done
