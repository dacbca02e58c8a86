#!/bin/bash
# N GPUs: symmetric heap (VMM + multicast), everything that now allocates from it, NVLS collectives
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29551 tests/mp/heap_check.py > gpurun_out/heap_check_$N.log 2>&1; echo "heap rc=$?"
grep -E "rank|PASSED|FAILED|Error|error|stats" gpurun_out/heap_check_$N.log | tail -14
timeout 300 $TR --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp${N}_check.log 2>&1; echo "stencil mp rc=$?"
grep -E "PASSED|FAILED|FAIL|Error" gpurun_out/mp${N}_check.log | tail -4
timeout 300 $TR --master-port 29541 tests/mp/coll_root_stress.py > gpurun_out/coll_root_stress_$N.log 2>&1; echo "root stress rc=$?"
grep -E "PASSED|FAILED|Error" gpurun_out/coll_root_stress_$N.log | tail -3
timeout 300 $TR --master-port 29542 tests/mp/dot_p2p_mp_check.py > gpurun_out/dot_p2p_mp_$N.log 2>&1; echo "dot/p2p rc=$?"
grep -E "PASSED|FAILED|FAIL|Error" gpurun_out/dot_p2p_mp_$N.log | tail -3
timeout 500 $TR --master-port 29620 bench/collectives.py --max-log2 ${MAXLOG2:-26} --json gpurun_out/collectives_${N}gpu.json > gpurun_out/coll_${N}.log 2>&1; echo "collectives rc=$?"
grep -E "^\{|PASSED|FAILED|Error|error" gpurun_out/coll_${N}.log | cut -c1-260 | tail -22
