# Convenience targets; everything they do is a plain command you can also type.
PY ?= python

.PHONY: build test test-gpu bench bench8 cmake clean

build:            ## compile csrc/ for sm_100a into cuda_mpi_scratch_b200/lib/libb200comm.so
	$(PY) -m cuda_mpi_scratch_b200.build

test: build       ## CPU suite (multi-rank logic on threads / processes, no GPU needed)
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: build   ## GPU suite (one B200)
	$(PY) -m pytest tests -x -q -m gpu

bench: build      ## headline benchmark on one GPU
	$(PY) bench.py --gpus 1 --steps 20 --warmup 3

bench8: build     ## headline benchmark on 8 GPUs of one node
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29500 \
	    bench.py --gpus 8 --steps 20 --warmup 3

cmake:            ## the same library + C++ examples through CMake
	cmake -S . -B build/cmake -G Ninja && cmake --build build/cmake

clean:
	rm -rf build cuda_mpi_scratch_b200/lib/libb200comm.so cuda_mpi_scratch_b200/lib/libb200comm.so.digest
