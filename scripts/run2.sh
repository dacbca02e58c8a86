# 2-GPU validation + benches (run with: gpurun --gpus 2 -- bash scripts/run2.sh)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29533 tests/mp/dot_p2p_mp_check.py > gpurun_out/mp2_dotp2p.log 2>&1; echo "rc=$?" >> gpurun_out/mp2_dotp2p.log
grep "OK\|FAIL\|PASSED\|Error\|error" gpurun_out/mp2_dotp2p.log | head -20
timeout 600 $TR --master-port 29600 bench/pingpong.py --json gpurun_out/pingpong2.json > gpurun_out/pingpong2.log 2>&1; echo "rc=$?" >> gpurun_out/pingpong2.log
tail -n 5 gpurun_out/pingpong2.log
timeout 300 $TR --master-port 29610 bench/dot.py --json gpurun_out/dot2.json > gpurun_out/dot2.log 2>&1; echo "rc=$?" >> gpurun_out/dot2.log
grep "impl\|rc=\|Error" gpurun_out/dot2.log | cut -c1-300
timeout 300 $TR --master-port 29620 bench/datatype_halo.py --json gpurun_out/datatype2.json > gpurun_out/datatype2.log 2>&1; echo "rc=$?" >> gpurun_out/datatype2.log
grep "case\|rc=\|Error" gpurun_out/datatype2.log | cut -c1-300
