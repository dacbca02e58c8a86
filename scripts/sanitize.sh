mkdir -p gpurun_out
timeout 200 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_default.json')); print(d['value'], d['ms_per_step'], d['clocks'], d['gpu_launches'], d['e2e']['value'])"
timeout 200 python bench.py --impl reference
timeout 600 python -m pytest tests/test_stencil_gpu.py tests/test_dot_gpu.py -m gpu -x -q -k "both_kernels or ones_give or residual" > gpurun_out/plain_sel.log 2>&1 && timeout 800 compute-sanitizer --tool memcheck --error-exitcode 3 python -m pytest tests/test_stencil_gpu.py tests/test_dot_gpu.py -m gpu -x -q -k "both_kernels or ones_give or residual" > gpurun_out/memcheck.log 2>&1; echo "memcheck rc=$?"; tail -n 6 gpurun_out/memcheck.log
