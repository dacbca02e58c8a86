#!/bin/bash
# PBS job: GPU ranks of the native drivers, one process per rank, wrapped in `time`.
# Counterpart of the reference's mpi_cuda_pbs_ref.sh:1-18 (4 nodes x 2 processes on Fermi
# GPUs under mpiexec.hydra, binary wrapped in `time`); the built-in launcher is single-node,
# so this asks for one node with 8 GPUs and starts 8 ranks (rank % devices picks the GPU).
#PBS -N b200-gpu-ranks
#PBS -l select=1:ncpus=16:ngpus=8
#PBS -l walltime=00:05:00
#PBS -j oe
set -euo pipefail
cd "${PBS_O_WORKDIR:-.}"
python -m cuda_mpi_scratch_b200.build
LIB=cuda_mpi_scratch_b200/lib
g++ -std=c++17 -O2 -Iinclude examples/cpp/dot_gpu.cpp -o dot_gpu -L$LIB -lb200comm -Wl,-rpath,$PWD/$LIB -pthread
time python -m cuda_mpi_scratch_b200.launch -n 8 ./dot_gpu 30 --variant 2
g++ -std=c++17 -O2 -Iinclude examples/cpp/pingpong_gpu.cpp -o pingpong_gpu -L$LIB -lb200comm -Wl,-rpath,$PWD/$LIB -pthread
time python -m cuda_mpi_scratch_b200.launch -n 2 ./pingpong_gpu 134217728 --async
