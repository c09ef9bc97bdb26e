#!/bin/bash
# Round-2 call n (N GPUs): 1 KiB - 1 GiB all-reduce sweep and the deep / wide MLP configs.
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=5000
N=${1:-8}; WHAT=${2:-sweep,deep,wide}
if [ "$N" = "1" ]; then RUN="python"; else RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612"; fi
if [[ $WHAT == *sweep* ]]; then
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench/allreduce_sweep.py --min-kb 1 --max-mb 1024 --iters 6 --out gpurun_out/sweep_n$N.json > gpurun_out/n_sweep_n$N.log 2>&1; echo "sweep rc=$?"; tail -13 gpurun_out/n_sweep_n$N.log | cut -c1-330
fi
if [[ $WHAT == *deep* ]]; then
timeout 600 $RUN bench/mlp_configs.py --config deep --steps 10 --warmup 3 --t1-ms ${T1_DEEP:-0} --out gpurun_out/deep_n$N.json > gpurun_out/n_deep_n$N.log 2>&1; echo "deep rc=$?"; tail -1 gpurun_out/n_deep_n$N.log | cut -c1-1500
fi
if [[ $WHAT == *wide* ]]; then
timeout 900 $RUN bench/mlp_configs.py --config wide --steps 6 --warmup 2 --t1-ms ${T1_WIDE:-0} --out gpurun_out/wide_n$N.json > gpurun_out/n_wide_n$N.log 2>&1; echo "wide rc=$?"; tail -1 gpurun_out/n_wide_n$N.log | cut -c1-1500
fi
