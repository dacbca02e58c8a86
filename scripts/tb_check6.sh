mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_stencil_gpu.py -m gpu -x -q > gpurun_out/tb6_tests.log 2>&1; echo "rc=$?" >> gpurun_out/tb6_tests.log; tail -n 4 gpurun_out/tb6_tests.log
run() { timeout 200 python bench.py --steps 48 --warmup 4 --temporal $1 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('T=$1', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'T1:', (d['one_time_step_per_sweep'] or {}).get('value'), d['roofline'])"; }
run 3
run 4
run 2
timeout 200 python bench.py --steps 48 --warmup 4 --temporal 3 --tile 4096 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('T=3 4096', d['value'], d['ms_per_step'])"
