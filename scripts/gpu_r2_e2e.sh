#!/bin/bash
# 1 GPU: stencil GPU tests (folded sweep tail, step_io), the end-to-end breakdown, default bench.py
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_stencil_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -8) > gpurun_out/pytest_stencil.log; cat gpurun_out/pytest_stencil.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 300 python bench/e2e_breakdown.py > gpurun_out/e2e_breakdown.json 2> gpurun_out/e2e_breakdown.err; echo "breakdown rc=$?"; cat gpurun_out/e2e_breakdown.json; tail -3 gpurun_out/e2e_breakdown.err
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_1gpu_io.json 2> gpurun_out/bench_1gpu_io.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=None
    for l in open("gpurun_out/bench_1gpu_io.json"):
        if l.startswith('{"metric"'): d=json.loads(l)
    print({k:d[k] for k in ("value","ms_per_step","ms_per_step_min","ms_per_step_max","vs_baseline","verified","gpu_launches")}, d["clocks"])
    print("   unfused", d["unfused_baseline"]["value"], "T1", d["one_time_step_per_sweep"]["value"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"])
except Exception as e:
    print("no json", e); print(open("gpurun_out/bench_1gpu_io.err").read()[-2000:])
PY
