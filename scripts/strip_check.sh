mkdir -p gpurun_out
for s in 3 4 5 6; do
B2C_STRIP_SLOTS=$s timeout 200 python bench.py --gpus 1 --steps 30 --warmup 3 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('slots=$s', d['value'], d['ms_per_step'], d['roofline']['fraction_of_measured_hbm'])"
done
for s in 4 5; do
B2C_STRIP_SLOTS=$s B2C_STRIP_CTAS_PER_SM=3 timeout 200 python bench.py --gpus 1 --steps 30 --warmup 3 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('slots=$s x3/sm', d['value'], d['ms_per_step'])"
done
B2C_STRIP_SLOTS=4 B2C_STRIP_CTAS_PER_SM=4 timeout 200 python bench.py --gpus 1 --steps 30 --warmup 3 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('slots=4 x4/sm', d['value'], d['ms_per_step'])"
