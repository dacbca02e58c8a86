mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log; tail -n 8 gpurun_out/pytest_gpu.log
timeout 240 $TR --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp2_check.log 2>&1; grep "MP_STENCIL\|FAIL\|graph\|Error" gpurun_out/mp2_check.log | head -6
timeout 200 $TR --master-port 29541 bench.py --gpus 2 --steps 50 --warmup 5 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('2gpu', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['gpu_launches'], d['clocks'])"
