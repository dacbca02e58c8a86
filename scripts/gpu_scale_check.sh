mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp${N}_check.log 2>&1; grep "MP_STENCIL\|FAIL\|Error" gpurun_out/mp${N}_check.log | head -6; grep -c " OK" gpurun_out/mp${N}_check.log
: > gpurun_out/scale_tb3_${N}.jsonl
timeout 200 $TR --master-port 29541 bench.py --gpus $N --steps 48 --warmup 4 2>gpurun_out/b${N}.err | grep metric >> gpurun_out/scale_tb3_${N}.jsonl
timeout 200 $TR --master-port 29543 bench.py --gpus $N --steps 48 --warmup 4 --scaling strong 2>/dev/null | grep metric >> gpurun_out/scale_tb3_${N}.jsonl
python - <<PY
import json
for l in open('gpurun_out/scale_tb3_${N}.jsonl'):
    d=json.loads(l); print(d['n_gpus'], d['scaling'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'T1', d['one_time_step_per_sweep']['value'], d['gpu_launches'], d['clocks']['reasons'], d['roofline']['fraction_of_measured_hbm'])
PY
