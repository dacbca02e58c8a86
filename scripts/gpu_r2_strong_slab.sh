#!/bin/bash
# N GPUs: strong scaling of 16384^2 with a row-slab decomposition (N x 1 process grid: tiles stay
# 16384 wide, the halo traffic is contiguous rows), next to the 1-GPU rate of the same box
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
show() {
python - $1 <<'PY'
import json,sys
try:
    d=None
    for l in open(sys.argv[1]):
        if l.startswith('{"metric"'): d=json.loads(l)
    print({k:d[k] for k in ("value","ms_per_step","ms_per_step_min","ms_per_step_max","vs_baseline","verified","steps")}, d["clocks"]["sm_mhz"], d["clocks"]["reasons"], d["config"]["tile_per_gpu"], d["config"]["process_grid"])
    u=d.get("unfused_baseline"); t1=d.get("one_time_step_per_sweep")
    print("   unfused", u and u["value"], "T1", t1 and t1["value"], "e2e", d["e2e"]["value"])
except Exception as e:
    print("no json", e); print(open(sys.argv[1].replace(".json",".err")).read()[-2000:])
PY
}
timeout 120 python bench.py --steps 20 --warmup 5 --no-baseline --min-seconds 1.0 > gpurun_out/bench_1gpu_samebox_slab.json 2> gpurun_out/bench_1gpu_samebox_slab.err; echo "1 GPU rc=$?"
show gpurun_out/bench_1gpu_samebox_slab.json
for K in 240 20; do
timeout 300 $TR --master-port 29536 bench.py --gpus $N --steps $K --warmup 5 --scaling strong --dims ${N}x1 --min-seconds 1.0 > gpurun_out/bench_${N}gpu_strong_slab_K$K.json 2> gpurun_out/bench_${N}gpu_strong_slab_K$K.err; echo "strong slab K=$K rc=$?"
show gpurun_out/bench_${N}gpu_strong_slab_K$K.json
done
