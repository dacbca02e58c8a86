#!/bin/bash
# temporal strip kernel: tests, bench (strip T=3, T=2), then one ncu capture
mkdir -p gpurun_out
(timeout 300 python -m pytest tests/test_stencil_gpu.py -x -q -k "temporal" 2>&1 | tail -5) > gpurun_out/r2c_pytest.log
cat gpurun_out/r2c_pytest.log
for cfg in ${CFGS:-strip:3 strip:2}; do
  set -- ${cfg/:/ }
  echo "== algo=$1 T=$2"
  B2C_TB_ALGO=$1 timeout 120 python bench.py --steps 24 --warmup 3 --temporal $2 --no-baseline --min-seconds 1.5 \
      > gpurun_out/r2c_bench_$1_T$2.json 2> gpurun_out/r2c_bench_$1_T$2.err
  echo "rc=$?"; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2c_bench_$1_T$2.json"))
    print({k:d[k] for k in ("value","ms_per_step","ms_per_step_min","ms_per_step_max","clocks","gpu_launches")}, d["e2e"]["value"], d["roofline"])
except Exception as e:
    print("no json", e); print(open("gpurun_out/r2c_bench_$1_T$2.err").read()[-2000:])
PY
done
bash scripts/ncu_tstrip.sh
