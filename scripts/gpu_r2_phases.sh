#!/bin/bash
# per-phase timing of the temporal strip kernel at 1 and 2 GPUs (same box), weak and strong
mkdir -p gpurun_out
export B2C_STENCIL_DEBUG=32
CUDA_VISIBLE_DEVICES=0 python bench/stencil_phases.py 2>/dev/null | grep '^{' > gpurun_out/phases_1gpu.jsonl
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$TR --master-port 29640 bench/stencil_phases.py 2>/dev/null | grep '^{' > gpurun_out/phases_2gpu.jsonl
$TR --master-port 29641 bench/stencil_phases.py --scaling strong 2>/dev/null | grep '^{' > gpurun_out/phases_2gpu_strong.jsonl
for f in gpurun_out/phases_*.jsonl; do echo $f; cut -c1-420 $f; done
