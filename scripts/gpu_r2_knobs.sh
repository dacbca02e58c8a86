#!/bin/bash
# temporal strip kernel: ring depth / strip width sweep (1 GPU)
mkdir -p gpurun_out
run() {
  echo "== $*"
  env "$@" timeout 120 python bench.py --steps 24 --warmup 3 --temporal 3 --no-baseline --min-seconds 1.0 > gpurun_out/knob.json 2> gpurun_out/knob.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/knob.json"))
    print(d["value"], d["ms_per_step"], d["ms_per_step_min"], d["clocks"]["sm_mhz"], d["clocks"]["power_w_max"], "e2e", d["e2e"]["value"])
except Exception as e:
    print("no json", e); print(open("gpurun_out/knob.err").read()[-1500:])
PY
}
run B2C_TS_SLOTS=2
run B2C_TS_SLOTS=3
run B2C_TS_SLOTS=4
run B2C_TS_SLOTS=3 B2C_TS_WIDTH=1536
run B2C_TS_SLOTS=3 B2C_TS_WIDTH=1176
run B2C_TS_SLOTS=2 B2C_TS_WIDTH=1536
run B2C_TS_SLOTS=3
