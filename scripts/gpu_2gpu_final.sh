mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp2_check.log 2>&1; grep "MP_STENCIL\|FAIL\|Error" gpurun_out/mp2_check.log | head -6; grep -c " OK" gpurun_out/mp2_check.log
timeout 200 $TR --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 3 2>/dev/null | grep metric > gpurun_out/bench_final_2gpu.json; python -c "import json; d=json.load(open('gpurun_out/bench_final_2gpu.json')); print(d['value'], d['steps'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])"
