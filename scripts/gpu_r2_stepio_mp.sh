#!/bin/bash
# N GPUs: multi-process stencil checks (incl. the pipelined step_io loop), the C++ stencil example, bench.py
N=${1:-2}
bash scripts/gpu_r2_mp.sh $N
(timeout 600 python -m pytest tests/test_cpp_examples.py -m gpu -x -q -k "stencil" 2>&1 | tail -4)
