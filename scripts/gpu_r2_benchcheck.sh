#!/bin/bash
# 1 GPU: default bench.py once, print the fields that matter
mkdir -p gpurun_out
timeout 80 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_1gpu_last.json 2> gpurun_out/bench_1gpu_last.err; echo "rc=$?"
wc -l gpurun_out/bench_1gpu_last.json
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench_1gpu_last.json").read())
    print(d["value"], d["e2e"]["value"], d["gpu_launches"], d["gpu_launches_per_batch"], d["timed_batches"],
          d["verified"], d["vs_baseline"], d["clocks"]["sm_mhz"])
except Exception as e:
    print("no json", e); print(open("gpurun_out/bench_1gpu_last.err").read()[-1500:])
PY
