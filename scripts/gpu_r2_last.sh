#!/bin/bash
# 1 GPU, end of round: default bench.py, smoke(), then the full -m gpu suite
mkdir -p gpurun_out
timeout 200 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_1gpu_final.json 2> gpurun_out/bench_1gpu_final.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=None
    for l in open("gpurun_out/bench_1gpu_final.json"):
        if l.startswith('{"metric"'): d=json.loads(l)
    print({k:d[k] for k in ("value","ms_per_step","vs_baseline","verified","gpu_launches")}, d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
    print("   unfused", d["unfused_baseline"]["value"], "T1", d["one_time_step_per_sweep"]["value"], "e2e", d["e2e"]["value"])
    print("   stdout lines:", sum(1 for _ in open("gpurun_out/bench_1gpu_final.json")))
except Exception as e:
    print("no json", e); print(open("gpurun_out/bench_1gpu_final.err").read()[-2000:])
PY
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
(timeout 280 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/pytest_gpu_final.log; cat gpurun_out/pytest_gpu_final.log
