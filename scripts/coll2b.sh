mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_collectives_gpu.py -m gpu -x -q 2>&1 | tail -3
bash scripts/coll2.sh 2
