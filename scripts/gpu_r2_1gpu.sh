#!/bin/bash
# 1 GPU: full -m gpu suite, smoke(), default bench.py, then ONE ncu capture of the temporal strip kernel
mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 200 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=None
    for l in open("gpurun_out/bench_1gpu.json"):
        if l.startswith('{"metric"'): d=json.loads(l)
    print({k:d[k] for k in ("value","ms_per_step","ms_per_step_min","ms_per_step_max","vs_baseline","verified","gpu_launches")}, d["clocks"])
    print("   unfused", d["unfused_baseline"]["value"], "T1", d["one_time_step_per_sweep"]["value"], "e2e", d["e2e"]["value"], d["roofline"])
except Exception as e:
    print("no json", e); print(open("gpurun_out/bench_1gpu.err").read()[-2000:])
PY
OUT=tstrip_T3_final bash scripts/ncu_tstrip.sh
