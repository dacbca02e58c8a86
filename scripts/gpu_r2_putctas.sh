#!/bin/bash
mkdir -p gpurun_out
for c in 16 32 96 148 296; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench/pingpong.py \
   --ctas $c --json gpurun_out/put_ctas_$c.json --paths kernel_tma,kernel_ldst > gpurun_out/put.log 2>&1
python - <<PY
import json
for r in json.load(open("gpurun_out/put_ctas_$c.json")):
    if r["bytes"] in (1<<21, 1<<24, 1<<27, 1<<30): print("ctas $c", r["path"], r["bytes"], r["rtt_us"], r["GBps_one_way"])
PY
done
