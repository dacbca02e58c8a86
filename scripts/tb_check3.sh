mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_stencil_gpu.py -m gpu -x -q -k temporal > gpurun_out/tb3_tests.log 2>&1; echo "rc=$?" >> gpurun_out/tb3_tests.log; tail -n 6 gpurun_out/tb3_tests.log
for T in 3 4; do for ST in 2 3 4; do
  B2C_TILE_STAGES=$ST timeout 200 python bench.py --steps 48 --warmup 4 --temporal $T 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('T=$T stages=$ST', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['roofline']['fraction_of_measured_hbm'])"
done; done
B2C_TILE_STAGES=3 timeout 200 python bench.py --steps 48 --warmup 4 --temporal 2 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('T=2 stages=3', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
