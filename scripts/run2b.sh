mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29620 bench/datatype_halo.py --json gpurun_out/datatype2.json > gpurun_out/datatype2.log 2>&1; echo "rc=$?" >> gpurun_out/datatype2.log
grep "case\|rc=\|Error" gpurun_out/datatype2.log | cut -c1-200
for ctas in 32 148 296; do
timeout 300 $TR --master-port 2960$((ctas%10)) bench/pingpong.py --ctas $ctas --paths kernel_ldst,kernel_tma,device_loop --json gpurun_out/pp_ctas$ctas.json > gpurun_out/pp_ctas$ctas.log 2>&1
echo "ctas=$ctas"; grep '"bytes": \(8\|2097152\|16777216\|134217728\|1073741824\),' gpurun_out/pp_ctas$ctas.log | cut -c1-140
done
timeout 120 $TR --master-port 29611 bench/dot.py --log2n 20 --iters 200 > gpurun_out/dot2_small.log 2>&1; grep "impl" gpurun_out/dot2_small.log | cut -c1-200
