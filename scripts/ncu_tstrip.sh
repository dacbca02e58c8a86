#!/bin/bash
# plain run first, then one ncu --set full capture of the temporal strip kernel (1 GPU)
mkdir -p gpurun_out
CMD="python bench.py --steps 6 --warmup 3 --temporal ${T:-3} --no-baseline --min-seconds 0.05"
B2C_TB_ALGO=${ALGO:-strip} $CMD > gpurun_out/ncu_plain.log 2>&1 &&
B2C_TB_ALGO=${ALGO:-strip} ncu --set full --clock-control none --import-source on -k regex:${KREGEX:-tstrip} -s 4 -c 2 \
    -f -o gpurun_out/${OUT:-tstrip_T3} $CMD > gpurun_out/ncu_run.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_run.log
