mkdir -p gpurun_out
timeout 200 python bench/profile_stencil.py 16384 6 1 --temporal 3 > gpurun_out/plain_tb3.log 2>&1; echo "plain rc=$?"; tail -1 gpurun_out/plain_tb3.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stencil_fused -c 2 -f -o gpurun_out/stencil_tb3 python bench/profile_stencil.py 16384 6 1 --temporal 3 > gpurun_out/ncu_tb3.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_tb3.log
