export B200COMM_NO_AUTOBUILD=1
run() { # T ctas stages
  B2C_TILE_STAGES=$2 timeout 200 python bench.py --steps 48 --warmup 4 --temporal $1 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('TH64 T=$1 stages=$2', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
}
timeout 300 python -m pytest tests/test_stencil_gpu.py -m gpu -x -q -k "temporal" 2>&1 | tail -2
run 3 2
run 2 2
run 4 2
run 1 2
