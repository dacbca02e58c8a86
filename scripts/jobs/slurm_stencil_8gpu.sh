#!/bin/bash
# SLURM job: fused stencil on one 8 x B200 node, one rank per GPU.
# Counterpart of the reference's batch scripts (mpi_cuda_pbs_ref.sh:1-18,
# stencil2d/sample-output/job_9_1_1_cuda-2d-stencil-subarray.slurm:1-15), which start
# `mpiexec.hydra -rmk slurm ./a.out`; here the launcher is torchrun (or the library's own
# `python -m cuda_mpi_scratch_b200.launch`), and the rank -> GPU binding comes from LOCAL_RANK.
#SBATCH --job-name=b200-stencil
#SBATCH --nodes=1
#SBATCH --ntasks-per-node=1
#SBATCH --gres=gpu:8
#SBATCH --time=00:10:00
#SBATCH --output=job_%j.out
set -euo pipefail
cd "${SLURM_SUBMIT_DIR:-.}"
python -m cuda_mpi_scratch_b200.build
time python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
    --master-port 29500 bench.py --gpus 8 --steps 100 --warmup 5
