mkdir -p gpurun_out
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 300 $TR8 --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp8_check.log 2>&1; grep "MP_STENCIL\|FAIL\|Error" gpurun_out/mp8_check.log | head -6
: > gpurun_out/scale_tb2.jsonl
timeout 200 $TR8 --master-port 29541 bench.py --gpus 8 --steps 40 --warmup 4 2>gpurun_out/b8.err | grep metric >> gpurun_out/scale_tb2.jsonl
timeout 200 $TR4 --master-port 29542 bench.py --gpus 4 --steps 40 --warmup 4 2>/dev/null | grep metric >> gpurun_out/scale_tb2.jsonl
timeout 200 $TR8 --master-port 29543 bench.py --gpus 8 --steps 40 --warmup 4 --scaling strong 2>/dev/null | grep metric >> gpurun_out/scale_tb2.jsonl
timeout 200 $TR4 --master-port 29544 bench.py --gpus 4 --steps 40 --warmup 4 --scaling strong 2>/dev/null | grep metric >> gpurun_out/scale_tb2.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/scale_tb2.jsonl'):
    d=json.loads(l); print(d['n_gpus'], d['scaling'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'T1', d['one_time_step_per_sweep'], d['gpu_launches'], d['clocks']['reasons'], d['roofline']['fraction_of_measured_hbm'])
PY
tail -3 gpurun_out/b8.err
