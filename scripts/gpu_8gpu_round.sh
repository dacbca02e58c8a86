# 8-GPU validation + scaling (gpurun --gpus 8 -- bash scripts/gpu_8gpu_round.sh)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu8.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu8.log; tail -n 3 gpurun_out/pytest_gpu8.log
timeout 300 $TR --nproc-per-node 8 --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp8_stencil.log 2>&1; echo "rc=$?" >> gpurun_out/mp8_stencil.log
grep "MP_STENCIL\|FAIL" gpurun_out/mp8_stencil.log | head -5
rm -f gpurun_out/scale.jsonl gpurun_out/scale_strong.jsonl
for n in 1 2 4 8; do
  if [ $n = 1 ]; then timeout 300 python bench.py --gpus 1 --steps 30 --warmup 3 2>/dev/null | grep metric >> gpurun_out/scale.jsonl
  else timeout 300 $TR --nproc-per-node $n --master-port 2954$n bench.py --gpus $n --steps 30 --warmup 3 2>/dev/null | grep metric >> gpurun_out/scale.jsonl; fi
done
for n in 2 4 8; do
  timeout 300 $TR --nproc-per-node $n --master-port 2955$n bench.py --gpus $n --steps 30 --warmup 3 --scaling strong 2>/dev/null | grep metric >> gpurun_out/scale_strong.jsonl
done
python - <<'PY'
import json
for f in ("gpurun_out/scale.jsonl","gpurun_out/scale_strong.jsonl"):
    base=None
    for l in open(f):
        d=json.loads(l)
        if base is None: base=d['value']/d['n_gpus']
        print(f, d['n_gpus'], d['scaling'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'per-gpu vs first', round(d['value']/d['n_gpus']/base,3))
PY
timeout 200 $TR --nproc-per-node 8 --master-port 29571 bench.py --gpus 8 --steps 30 --warmup 3 --radius 2 2>/dev/null | grep metric | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('8gpu radius2', d['value'], d['ms_per_step'])"
