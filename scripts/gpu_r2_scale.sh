#!/bin/bash
# N GPUs (4 or 8): heap, multi-process correctness, collectives vs NCCL (NVLS), weak + strong bench, dot
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29551 tests/mp/heap_check.py > gpurun_out/heap_check_$N.log 2>&1; echo "heap rc=$?"
grep -E "PASSED|FAILED|Error|stats" gpurun_out/heap_check_$N.log | tail -3
timeout 300 $TR --master-port 29533 tests/mp/stencil_mp_check.py > gpurun_out/mp${N}_check.log 2>&1; echo "stencil mp rc=$?"
grep -E "PASSED|FAILED|FAIL|Error" gpurun_out/mp${N}_check.log | tail -3
B2C_TB_ALGO=strip timeout 300 $TR --master-port 29534 tests/mp/stencil_mp_check.py > gpurun_out/mp${N}_check_strip.log 2>&1; echo "stencil mp (strip) rc=$?"
grep -E "PASSED|FAILED|FAIL|Error" gpurun_out/mp${N}_check_strip.log | tail -3
timeout 300 $TR --master-port 29541 tests/mp/coll_root_stress.py > gpurun_out/coll_root_stress_$N.log 2>&1; echo "root stress rc=$?"
grep -E "PASSED|FAILED|Error" gpurun_out/coll_root_stress_$N.log | tail -2
timeout 600 $TR --master-port 29620 bench/collectives.py --max-log2 ${MAXLOG2:-30} --json gpurun_out/collectives_${N}gpu.json > gpurun_out/coll_${N}.log 2>&1; echo "collectives rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/collectives_${N}gpu.json"))
    print("heap", d.get("symmetric_heap"), d["check"])
    for r in d["rows"]:
        print({k:v for k,v in r.items() if k!="clocks"})
except Exception as e:
    print("no json", e); print(open("gpurun_out/coll_${N}.log").read()[-3000:])
PY
show() { python - "$1" <<'PY'
import json,sys
try:
    d=None
    for l in open(sys.argv[1]):
        if l.startswith('{"metric"'): d=json.loads(l)
    print({k:d[k] for k in ("value","ms_per_step","ms_per_step_min","ms_per_step_max","vs_baseline","verified","scaling")}, d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
    print("   unfused", d["unfused_baseline"]["value"], "T1", d["one_time_step_per_sweep"]["value"], "e2e", d["e2e"]["value"], d["roofline"]["fraction_of_measured_hbm"], d["config"]["tile_per_gpu"])
except Exception as e:
    print("no json", e); print(open(sys.argv[1].replace(".json",".err")).read()[-2000:])
PY
}
timeout 300 $TR --master-port 29535 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench weak rc=$?"; show gpurun_out/bench_${N}gpu.json
timeout 300 $TR --master-port 29536 bench.py --gpus $N --steps 20 --warmup 5 --scaling strong > gpurun_out/bench_${N}gpu_strong.json 2> gpurun_out/bench_${N}gpu_strong.err; echo "bench strong rc=$?"; show gpurun_out/bench_${N}gpu_strong.json
B2C_TB_ALGO=strip timeout 300 $TR --master-port 29537 bench.py --gpus $N --steps 20 --warmup 5 --scaling strong --no-baseline > gpurun_out/bench_${N}gpu_strong_strip.json 2> gpurun_out/bench_${N}gpu_strong_strip.err; echo "bench strong (strip forced) rc=$?"
grep -o '"value": [0-9.]*, "unit": "Gcell/s", "n_gpus"' gpurun_out/bench_${N}gpu_strong_strip.json | head -1
timeout 300 $TR --master-port 29538 bench/dot.py --json gpurun_out/dot_${N}gpu.json > gpurun_out/dot_${N}gpu.log 2>&1; echo "dot rc=$?"
grep '^{' gpurun_out/dot_${N}gpu.log | cut -c1-330
