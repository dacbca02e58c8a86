# temporal blocking + native GPU drivers: tests, then 1-GPU bench with and without --temporal 2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_stencil_gpu.py tests/test_cpp_examples.py -m gpu -x -q > gpurun_out/tb_tests.log 2>&1; echo "rc=$?" >> gpurun_out/tb_tests.log; tail -n 12 gpurun_out/tb_tests.log
timeout 60 python examples/python/dot_single_gpu.py 1024 > gpurun_out/dot_single.log 2>&1; tail -n 7 gpurun_out/dot_single.log
for T in 1 2; do
  timeout 200 python bench.py --steps 40 --warmup 4 --temporal $T 2>gpurun_out/tb_bench_$T.err | grep metric > gpurun_out/tb_bench_$T.json
  python -c "import sys,json; d=json.loads(open('gpurun_out/tb_bench_$T.json').read()); print('T=$T', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['gpu_launches'], d['roofline'])" || tail -5 gpurun_out/tb_bench_$T.err
done
for ST in 2 3; do
  B2C_TILE_STAGES=$ST timeout 200 python bench.py --steps 40 --warmup 4 --temporal 2 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('T=2 stages=$ST', d['value'], d['ms_per_step'])"
done
timeout 200 python bench.py --steps 40 --warmup 4 --temporal 2 --tile 4096 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('T=2 4096', d['value'], d['ms_per_step'])"
timeout 200 python bench.py --steps 40 --warmup 4 --temporal 1 --tile 4096 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('T=1 4096', d['value'], d['ms_per_step'])"
