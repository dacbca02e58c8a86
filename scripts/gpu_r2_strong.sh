#!/bin/bash
# N GPUs: strong scaling with batches long enough (K = 240 steps) to amortise the start skew between ranks
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for K in 240 20; do
timeout 300 $TR --master-port 29536 bench.py --gpus $N --steps $K --warmup 5 --scaling strong --min-seconds 1.0 > gpurun_out/bench_${N}gpu_strong_K$K.json 2> gpurun_out/bench_${N}gpu_strong_K$K.err; echo "strong K=$K rc=$?"
python - gpurun_out/bench_${N}gpu_strong_K$K.json <<'PY'
import json,sys
try:
    d=None
    for l in open(sys.argv[1]):
        if l.startswith('{"metric"'): d=json.loads(l)
    print({k:d[k] for k in ("value","ms_per_step","ms_per_step_min","ms_per_step_max","vs_baseline","verified","steps")}, d["clocks"]["sm_mhz"], d["clocks"]["reasons"], d["config"]["tile_per_gpu"], d["config"]["kernel"][:24])
    print("   unfused", d["unfused_baseline"]["value"], "T1", d["one_time_step_per_sweep"]["value"], "e2e", d["e2e"]["value"])
except Exception as e:
    print("no json", e); print(open(sys.argv[1].replace(".json",".err")).read()[-2000:])
PY
done
