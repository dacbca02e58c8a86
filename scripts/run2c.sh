mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp2_check.log 2>&1; echo "rc=$?" >> gpurun_out/mp2_check.log
grep "MP_STENCIL\|FAIL" gpurun_out/mp2_check.log | head
for algo in strip tile; do
B2C_STENCIL_ALGO=$algo timeout 200 $TR --master-port 29541 bench.py --gpus 2 --steps 30 --warmup 3 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('2gpu $algo', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
done
timeout 200 python bench.py --gpus 1 --steps 30 --warmup 3 > gpurun_out/bench1_strip.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/bench1_strip.json')); print('1gpu', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
