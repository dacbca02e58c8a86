mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_stencil_gpu.py -m gpu -x -q -k both_kernels 2>&1 | tail -n 2
for s in 6 8 10 12; do
B2C_STENCIL_ALGO=strip B2C_STRIP_SLOTS=$s timeout 200 python bench.py --gpus 1 --steps 30 --warmup 3 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('chunk2 slots=$s', d['value'], d['ms_per_step'], d['roofline']['fraction_of_measured_hbm'])"
done
for s in 6 8; do
B2C_STENCIL_ALGO=strip B2C_STRIP_SLOTS=$s B2C_STRIP_CTAS_PER_SM=3 timeout 200 python bench.py --gpus 1 --steps 30 --warmup 3 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('chunk2 slots=$s x3/sm', d['value'], d['ms_per_step'])"
done
