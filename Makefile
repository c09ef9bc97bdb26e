# b200-ainic build / test / bench shortcuts (counterpart of the reference's hw/Makefile + sw/Makefile).
PY ?= python

.PHONY: all build force sass test test-gpu bench bench-ddp sweep clean

all: build

build:            ## compile libainic_cuda.so (sm_100a), libainic_oracle.so and the CPU app
	$(PY) -m fpga_ai_nic_b200.utils.build

force:
	$(PY) -m fpga_ai_nic_b200.utils.build --force

sass: build       ## SASS listing of every kernel -> profiles/sass/
	$(PY) -m fpga_ai_nic_b200.utils.build --sass

test: build       ## CPU tier (oracle, fake NIC, ring model, CPU app)
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: build   ## GPU tier (needs a B200)
	$(PY) -m pytest tests -x -q -m gpu

bench: build      ## headline: MLP training samples/s on 1 GPU
	$(PY) bench.py

bench-ddp:
	$(PY) bench.py --impl ddp

sweep: build      ## all-reduce + weight-update bandwidth sweep on all visible GPUs
	$(PY) -m fpga_ai_nic_b200.parallel.launcher --gpus $$(nvidia-smi -L | wc -l) -- bench/allreduce_sweep.py --max-mb 1024 --out gpurun_out/sweep.json

clean:
	rm -rf build fpga_ai_nic_b200/_lib/*.so
