mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_final.log; tail -n 5 gpurun_out/pytest_gpu_final.log
timeout 300 python -c "import __graft_entry__ as g; g.build(); g.smoke()" 2>&1 | tail -3
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 2>gpurun_out/bench_final.err | grep metric > gpurun_out/bench_final_1gpu.json; python -c "import json; d=json.load(open('gpurun_out/bench_final_1gpu.json')); print(d['value'], d['steps'], d['ms_per_step'], d['e2e'], d['gpu_launches'], d['clocks'], d['one_time_step_per_sweep'])"
