#!/bin/bash
# N GPUs: in-place bcast / allgather on symmetric tensors: example (correctness) + size sweep
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29611 examples/python/allreduce_in_switch.py 2>&1 | grep -E "PASSED|FAILED|Error|error" | head -5
bash scripts/gpu_r2_moves.sh $N
