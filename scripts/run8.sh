# 8-GPU validation + scaling (gpurun --gpus 8 -- bash scripts/run8.sh)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp8_stencil.log 2>&1; echo "rc=$?" >> gpurun_out/mp8_stencil.log
grep "MP_STENCIL\|FAIL" gpurun_out/mp8_stencil.log | head -5
timeout 300 $TR --nproc-per-node 8 --master-port 29534 tests/mp/dot_p2p_mp_check.py > gpurun_out/mp8_dotp2p.log 2>&1; echo "rc=$?" >> gpurun_out/mp8_dotp2p.log
grep "MP_DOT\|FAIL" gpurun_out/mp8_dotp2p.log | head -5
rm -f gpurun_out/scale.jsonl
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
for n in 4 8; do timeout 300 $TR --nproc-per-node $n --master-port 2961$n bench/dot.py --json gpurun_out/dot$n.json 2>/dev/null | grep impl | cut -c1-160; done
