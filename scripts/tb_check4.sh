mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_stencil_gpu.py -m gpu -x -q > gpurun_out/tb4_tests.log 2>&1; echo "rc=$?" >> gpurun_out/tb4_tests.log; tail -n 4 gpurun_out/tb4_tests.log
run() { # T ctas stages
  B2C_TILE_CTAS_PER_SM=$2 B2C_TILE_STAGES=$3 timeout 200 python bench.py --steps 48 --warmup 4 --temporal $1 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('T=$1 ctas=$2 stages=$3', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'T1:', (d['one_time_step_per_sweep'] or {}).get('value'))"
}
run 2 1 3
run 2 2 2
run 3 1 3
run 3 2 2
run 4 2 2
run 1 1 2
