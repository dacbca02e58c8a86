#!/bin/bash
# N GPUs: multi-process stencil check (default kernels, then temporal strip kernel forced), bench.py
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp${N}_check.log 2>&1; echo "mp check rc=$?"
grep -E "PASSED|FAILED|FAIL" gpurun_out/mp${N}_check.log | tail -5
B2C_TB_ALGO=strip timeout 300 $TR --master-port 29534 tests/mp/stencil_mp_check.py > gpurun_out/mp${N}_check_strip.log 2>&1; echo "mp check (strip) rc=$?"
grep -E "PASSED|FAILED|FAIL" gpurun_out/mp${N}_check_strip.log | tail -5
timeout 300 $TR --master-port 29535 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_${N}gpu.json"))
    print({k:d[k] for k in ("value","ms_per_step","ms_per_step_min","ms_per_step_max","vs_baseline","verified","clocks","gpu_launches")})
    print("unfused", d["unfused_baseline"]["value"], "T1", d["one_time_step_per_sweep"]["value"], "e2e", d["e2e"]["value"], d["roofline"])
except Exception as e:
    print("no json", e); print(open("gpurun_out/bench_${N}gpu.err").read()[-3000:])
PY
if [ -n "$STRONG" ]; then
timeout 300 $TR --master-port 29536 bench.py --gpus $N --steps 20 --warmup 5 --scaling strong > gpurun_out/bench_${N}gpu_strong.json 2> gpurun_out/bench_${N}gpu_strong.err; echo "strong rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_${N}gpu_strong.json"))
    print("strong", {k:d[k] for k in ("value","ms_per_step","vs_baseline","verified")}, "unfused", d["unfused_baseline"]["value"], "T1", d["one_time_step_per_sweep"]["value"])
except Exception as e:
    print("no json", e); print(open("gpurun_out/bench_${N}gpu_strong.err").read()[-3000:])
PY
fi
