#!/bin/bash
# N GPUs: full -m gpu suite (virtual ranks on GPU 0), then the multi-process checks of the
# collectives / dot / comm layer incl. the slow-root stress test
N=${1:-2}
mkdir -p gpurun_out
(timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/coll_pytest.log; cat gpurun_out/coll_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29541 tests/mp/coll_root_stress.py > gpurun_out/coll_root_stress_$N.log 2>&1; echo "root stress rc=$?"
grep -E "PASSED|FAILED|mismatches|Error|error" gpurun_out/coll_root_stress_$N.log | tail -6
timeout 300 $TR --master-port 29542 tests/mp/dot_p2p_mp_check.py > gpurun_out/dot_p2p_mp_$N.log 2>&1; echo "dot/p2p/comm rc=$?"
grep -E "PASSED|FAILED|FAIL|Error" gpurun_out/dot_p2p_mp_$N.log | tail -6
