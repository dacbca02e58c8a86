mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29620 bench/collectives.py --json gpurun_out/collectives_${N}gpu.json > gpurun_out/coll_${N}.log 2>&1; echo "rc=$?" >> gpurun_out/coll_${N}.log
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/coll_${N}.log | tail -25
