#!/bin/bash
# 2+ GPUs: the new examples / wrappers (C++ DeviceCollectives, collectives over CommWorld and TorchWorld)
N=${1:-2}
mkdir -p gpurun_out
g++ -std=c++17 -O1 -Iinclude examples/cpp/collectives_gpu.cpp -o gpurun_out/collectives_gpu -Lcuda_mpi_scratch_b200/lib -lb200comm -Wl,-rpath,$PWD/cuda_mpi_scratch_b200/lib -pthread
env -u WORLD_SIZE -u RANK -u LOCAL_RANK python -m cuda_mpi_scratch_b200.launch -n $N --timeout 120 gpurun_out/collectives_gpu 2>&1 | tail -3
env -u WORLD_SIZE -u RANK -u LOCAL_RANK python -m cuda_mpi_scratch_b200.launch -n $N --timeout 120 examples/python/allreduce_in_switch.py 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29650 examples/python/allreduce_in_switch.py 2>&1 | grep -E "PASSED|FAILED|Error" | tail -3
(CUDA_VISIBLE_DEVICES=0 timeout 300 python -m pytest tests/test_cpp_examples.py tests/test_heap_gpu.py -q -m gpu 2>&1 | tail -3)
