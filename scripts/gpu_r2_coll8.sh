#!/bin/bash
# N GPUs: collectives after the single-launch small-message path: stress + bench (small sizes matter)
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29541 tests/mp/coll_root_stress.py > gpurun_out/coll_root_stress_$N.log 2>&1; echo "root stress rc=$?"
grep -E "PASSED|FAILED|Error|mismatch" gpurun_out/coll_root_stress_$N.log | tail -4
timeout 600 $TR --master-port 29620 bench/collectives.py --max-log2 ${MAXLOG2:-22} --json gpurun_out/collectives_${N}gpu_small.json > gpurun_out/coll_${N}.log 2>&1; echo "collectives rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/collectives_${N}gpu_small.json"))
    print(d["check"])
    for r in d["rows"]:
        print({k:v for k,v in r.items() if k!="clocks"})
except Exception as e:
    print("no json", e); print(open("gpurun_out/coll_${N}.log").read()[-3000:])
PY
