mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_stencil_gpu.py -m gpu -x -q 2>&1 | tail -n 2
for shape in "16384 16384" "16384 8192" "8192 8192" "8192 4096" "4096 4096" "2048 2048"; do
set -- $shape
for algo in strip tile; do
B2C_STENCIL_ALGO=$algo timeout 200 python bench.py --gpus 1 --steps 50 --warmup 5 --th $1 --tw $2 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('tile $1x$2 $algo', d['value'], d['ms_per_step'])"
done
done
