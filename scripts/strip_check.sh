mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_stencil_gpu.py -m gpu -x -q 2>&1 | tail -n 2
for per in 2 3 4; do
B2C_STRIP_CTAS_PER_SM=$per B2C_STENCIL_ALGO=strip timeout 200 python bench.py --gpus 1 --steps 30 --warmup 3 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ctas/sm=$per', d['value'], d['ms_per_step'], d['roofline']['fraction_of_measured_hbm'])"
done
for w in 252 128; do
B2C_STRIP_WIDTH=$w B2C_STRIP_CTAS_PER_SM=2 timeout 200 python bench.py --gpus 1 --steps 30 --warmup 3 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('width=$w', d['value'], d['ms_per_step'])"
done
