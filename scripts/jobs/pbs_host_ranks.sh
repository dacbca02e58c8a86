#!/bin/bash
# PBS job: host-only multi-rank run over the shared-memory fabric (no GPUs, no MPI).
# Counterpart of mpi_pbs_sample.sh:1-18 (4 nodes x 16 MPI processes under mpiexec.hydra);
# the built-in transport is single-node, so this asks for one node and 16 processes.
#PBS -N b200-host-ranks
#PBS -l select=1:ncpus=16
#PBS -l walltime=00:00:30
#PBS -j oe
set -euo pipefail
cd "${PBS_O_WORKDIR:-.}"
python -m cuda_mpi_scratch_b200.launch -n 16 examples/python/ring_processes.py
