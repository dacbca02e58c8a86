mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 python -m pytest tests/test_stencil_gpu.py -m gpu -x -q 2>&1 | tail -n 2
timeout 240 $TR --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp2_check.log 2>&1; grep "MP_STENCIL\|FAIL" gpurun_out/mp2_check.log | head -3
for algo in strip tile; do
for rep in 1 2; do
B2C_STENCIL_ALGO=$algo timeout 200 $TR --master-port 2954$rep bench.py --gpus 2 --steps 50 --warmup 5 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('2gpu $algo', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
done
done
for algo in strip tile; do
B2C_STENCIL_ALGO=$algo timeout 200 python bench.py --gpus 1 --steps 50 --warmup 5 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('1gpu $algo', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
done
