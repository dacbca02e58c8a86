#!/bin/bash
# 2 GPUs: NVLink peer peak of this box, one direction and both at once (put kernels, copy engine, NCCL)
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench/pingpong.py \
   --json gpurun_out/nvlink_peak_2gpu.json --paths kernel_tma,kernel_ldst,bidir_tma,bidir_ldst,copy_engine,nccl > gpurun_out/nvlink.log 2>&1; echo rc=$?
python - <<'PY'
import json
for r in json.load(open("gpurun_out/nvlink_peak_2gpu.json")):
    if r["bytes"] >= 1<<27: print(r["path"], r["bytes"], r["rtt_us"], r["GBps_one_way"], r["check"], r["clocks"].get("sm_mhz"))
PY
