#!/bin/bash
# Round-2 call h (N GPUs): which of this round's changes slowed the large-bucket sweep (prefetch depth, carve-out)?
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AINIC_TIMEOUT_MS=3000
N=${1:-2}
for cfg in "8 1" "0 1" "8 0" "0 0" "3 1"; do set -- $cfg; 
AINIC_PF_DEPTH=$1 AINIC_COMM_CARVEOUT=$2 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29573 bench/allreduce_sweep.py --min-kb 65536 --max-mb 256 --codecs bfp,f32 --algos oneshot,twoshot --iters 6 --out gpurun_out/h_sweep_pf$1_c$2.json > gpurun_out/h_sweep_pf$1_c$2.log 2>&1; echo "pf=$1 carve=$2 rc=$?"; tail -2 gpurun_out/h_sweep_pf$1_c$2.log | cut -c1-200
done
