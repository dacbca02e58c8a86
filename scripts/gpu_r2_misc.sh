#!/bin/bash
# 2 GPUs: datatype halo bench (clocks, nested multi-buffer type); then on GPU 0 alone the ncu
# capture of the strided-column move
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29620 bench/datatype_halo.py --json gpurun_out/datatype_halo_2gpu.json > gpurun_out/datatype_halo_2gpu.log 2>&1; echo "datatype rc=$?"
grep '^{' gpurun_out/datatype_halo_2gpu.log | cut -c1-250
CUDA_VISIBLE_DEVICES=0 python bench/profile_column.py > gpurun_out/col_plain.log 2>&1 &&
CUDA_VISIBLE_DEVICES=0 ncu --set full --clock-control none --import-source on -k regex:uniform_move -s 3 -c 1 -f -o gpurun_out/column_g1 python bench/profile_column.py > gpurun_out/col_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/col_ncu.log
