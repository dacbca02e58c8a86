# 2 GPUs: multi-process correctness (incl. temporal blocking), C++ dot driver with the real
# one-launch fused path, then weak-scaling bench with --temporal 2 (ring depth 3/4/5)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp2_check.log 2>&1; grep "MP_STENCIL\|FAIL\|graph\|Error\|temporal 2" gpurun_out/mp2_check.log | head -12
timeout 300 python -m pytest tests/test_cpp_examples.py -m gpu -x -q -k "dot or pingpong" > gpurun_out/cpp2.log 2>&1; tail -n 3 gpurun_out/cpp2.log
for ST in 3 4 5; do
  B2C_TILE_STAGES=$ST timeout 200 python bench.py --steps 40 --warmup 4 --temporal 2 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('1gpu T=2 stages=$ST', d['value'], d['ms_per_step'])"
done
for ST in 3 4; do
  B2C_TILE_STAGES=$ST timeout 200 $TR --master-port 2954$ST bench.py --gpus 2 --steps 40 --warmup 4 --temporal 2 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('2gpu T=2 stages=$ST', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['gpu_launches'], d['clocks'])"
done
timeout 200 $TR --master-port 29547 bench.py --gpus 2 --steps 40 --warmup 4 --temporal 2 --scaling strong 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('2gpu strong T=2', d['value'], d['ms_per_step'])"
