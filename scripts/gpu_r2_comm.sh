#!/bin/bash
# 2 GPUs: comm-layer GPU tests (eager mailbox path), reference CUDA programs over the shim,
# multi-process comm check, ping-pong bench (comm API vs NCCL vs copy engine vs put kernel)
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_p2p_gpu.py tests/test_reference_conformance_gpu.py -x -q 2>&1 | tail -15) > gpurun_out/comm_pytest.log; cat gpurun_out/comm_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29542 tests/mp/dot_p2p_mp_check.py > gpurun_out/dot_p2p_mp_2.log 2>&1; echo "dot/p2p/comm rc=$?"
grep -E "PASSED|FAILED|FAIL|Error" gpurun_out/dot_p2p_mp_2.log | tail -6
timeout 400 $TR --master-port 29600 bench/pingpong.py --json gpurun_out/pingpong_2gpu.json \
   --paths ${PATHS:-kernel_tma,device_loop,copy_engine,nccl,comm_sync,comm_async} > gpurun_out/pingpong_2gpu.log 2>&1; echo "pingpong rc=$?"
grep device_sends gpurun_out/pingpong_2gpu.log
g++ -std=c++17 -O1 -Iinclude examples/cpp/pingpong_gpu.cpp -o gpurun_out/pingpong_gpu -Lcuda_mpi_scratch_b200/lib -lb200comm -Wl,-rpath,$PWD/cuda_mpi_scratch_b200/lib -pthread
for n in 1 512 4096 8192 131072; do
  python -m cuda_mpi_scratch_b200.launch -n 2 --timeout 120 gpurun_out/pingpong_gpu $n --iters 200 2>&1 | grep -v PASSED
  python -m cuda_mpi_scratch_b200.launch -n 2 --timeout 120 gpurun_out/pingpong_gpu $n --iters 200 --async 2>&1 | grep "round-trip"
done
B200COMM_NO_EAGER=1 python -m cuda_mpi_scratch_b200.launch -n 2 --timeout 120 gpurun_out/pingpong_gpu 1 --iters 200 2>&1 | grep "round-trip"
python - <<'PY'
import json
try:
    for r in json.load(open("gpurun_out/pingpong_2gpu.json")):
        if r["bytes"] in (8, 4096, 32768, 1<<21, 1<<30):
            print(r["path"], r["bytes"], r["rtt_us"], r["GBps_one_way"], r["check"], r["clocks"].get("sm_mhz"), r["clocks"].get("reasons"))
except Exception as e:
    print("no json", e); print(open("gpurun_out/pingpong_2gpu.log").read()[-3000:])
PY
