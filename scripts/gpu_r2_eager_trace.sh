#!/bin/bash
mkdir -p gpurun_out
g++ -std=c++17 -O1 -Iinclude examples/cpp/pingpong_gpu.cpp -o gpurun_out/pingpong_gpu -Lcuda_mpi_scratch_b200/lib -lb200comm -Wl,-rpath,$PWD/cuda_mpi_scratch_b200/lib -pthread
B200COMM_EAGER_TRACE=1 python -m cuda_mpi_scratch_b200.launch -n 2 --timeout 120 gpurun_out/pingpong_gpu 1 --iters 12 > gpurun_out/eager_trace.log 2>&1
tail -70 gpurun_out/eager_trace.log
