#!/bin/bash
# 2-GPU box: the checks the driver runs on GPU 0 (full -m gpu suite, smoke, bench), the CommWorld
# check on two GPUs, the 2-GPU bench, then ONE ncu capture (shared-memory conflicts of the strip kernel)
mkdir -p gpurun_out
(CUDA_VISIBLE_DEVICES=0 timeout 700 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
CUDA_VISIBLE_DEVICES=0 timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
env -u WORLD_SIZE -u RANK -u LOCAL_RANK timeout 200 python -m cuda_mpi_scratch_b200.launch -n 2 --timeout 180 tests/mp/commworld_check.py 2>&1 | grep -E "rank|PASSED|FAILED|Error" | tail -6
show() { python - "$1" <<'PY'
import json,sys
try:
    d=None
    for l in open(sys.argv[1]):
        if l.startswith('{"metric"'): d=json.loads(l)
    print({k:d[k] for k in ("value","ms_per_step","ms_per_step_min","ms_per_step_max","vs_baseline","verified","n_gpus")}, d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
    print("   unfused", d["unfused_baseline"]["value"], "T1", d["one_time_step_per_sweep"]["value"], "e2e", d["e2e"]["value"], d["roofline"]["fraction_of_measured_hbm"])
except Exception as e:
    print("no json", e); print(open(sys.argv[1].replace(".json",".err")).read()[-2000:])
PY
}
CUDA_VISIBLE_DEVICES=0 timeout 200 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench 1 rc=$?"; show gpurun_out/bench_1gpu.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench 2 rc=$?"; show gpurun_out/bench_2gpu.json
python bench.py --impl reference
CUDA_VISIBLE_DEVICES=0 OUT=tstrip_T3_final bash scripts/ncu_tstrip.sh | tail -2
