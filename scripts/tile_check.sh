mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_stencil_gpu.py -m gpu -x -q 2>&1 | tail -n 2
for shape in "16384 16384" "8192 4096" "4096 4096" "2048 2048"; do
for cfg in "2 2" "3 2"; do
set -- $shape $cfg
B2C_TILE_STAGES=$3 B2C_TILE_CTAS_PER_SM=$4 timeout 200 python bench.py --gpus 1 --steps 50 --warmup 5 --th $1 --tw $2 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('TH16 $1x$2 stages=$3 ctas/sm=$4', d['value'], d['ms_per_step'])"
done
done
