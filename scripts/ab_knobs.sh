mkdir -p gpurun_out; rm -f gpurun_out/ab2.log
for dbg in 0 4 8 12 24 28; do
  B2C_STENCIL_DEBUG=$dbg timeout 200 python bench.py --gpus 1 --steps 30 --warmup 3 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('debug=$dbg', d['value'], d['ms_per_step'], d['roofline']['fraction_of_measured_hbm'])" >> gpurun_out/ab2.log 2>&1
done
cat gpurun_out/ab2.log
