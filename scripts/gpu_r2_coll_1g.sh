#!/bin/bash
# N GPUs: the large end of the allreduce table (256 MiB and 1 GiB) against NCCL
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 500 $TR --master-port 29620 bench/collectives.py --min-log2 28 --max-log2 30 --json gpurun_out/collectives_${N}gpu_1g.json > gpurun_out/coll_${N}_1g.log 2>&1; echo "collectives rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/collectives_${N}gpu_1g.json"))
    print(d["check"])
    for r in d["rows"]:
        print({k:v for k,v in r.items() if k!="clocks"})
except Exception as e:
    print("no json", e); print(open("gpurun_out/coll_${N}_1g.log").read()[-3000:])
PY
