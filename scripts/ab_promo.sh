mkdir -p gpurun_out; rm -f gpurun_out/ab.log
for promo in 0 64 128 256; do
  B2C_STENCIL_L2PROMO=$promo timeout 200 python bench.py --gpus 1 --steps 30 --warmup 3 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('promo=$promo', d['value'], d['ms_per_step'], d['roofline']['fraction_of_measured_hbm'])" >> gpurun_out/ab.log 2>&1
done
timeout 200 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 2 >> gpurun_out/ab.log
cat gpurun_out/ab.log
