# ncu capture of the temporally blocked stencil kernel (after a plain run of the same program exited 0)
mkdir -p gpurun_out
timeout 200 python bench/profile_stencil.py --temporal 2 > gpurun_out/plain_tb2.log 2>&1; echo "plain rc=$?"; tail -2 gpurun_out/plain_tb2.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stencil_fused -c 2 -f -o gpurun_out/stencil_tb2 python bench/profile_stencil.py --temporal 2 > gpurun_out/ncu_tb2.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_tb2.log
