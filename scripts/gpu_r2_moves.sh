#!/bin/bash
# N GPUs: bcast and allgather over message sizes against NCCL (allreduce sizes skipped)
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29620 bench/collectives.py --min-log2 31 --max-log2 28 --sweep-moves --json gpurun_out/collectives_${N}gpu_moves.json > gpurun_out/coll_${N}_moves.log 2>&1; echo "collectives rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/collectives_${N}gpu_moves.json"))
    print(d["check"])
    for r in d["rows"]:
        print({k:v for k,v in r.items() if k!="clocks"})
except Exception as e:
    print("no json", e); print(open("gpurun_out/coll_${N}_moves.log").read()[-3000:])
PY
