#!/bin/bash
# SASS listings of every kernel family (cuobjdump -sass of the built library), one file per kernel,
# plus a one-line summary of the instructions that prove the design (artefacts/SUMMARY.txt).
# Runs on the CPU box: nvcc cross-compiles, no GPU needed.
set -e
LIB=cuda_mpi_scratch_b200/lib/libb200comm.so
OUT=artefacts
mkdir -p $OUT
cuobjdump -sass $LIB | grep "Function :" | sed 's/.*Function : //' > /tmp/b2c_functions.txt
dump() {   # name regex
  local fn
  fn=$(grep -E "$2" /tmp/b2c_functions.txt | head -1)
  if [ -z "$fn" ]; then echo "no kernel matches $2" >&2; return; fi
  cuobjdump -sass -fun "$fn" $LIB | grep -v "^Fatbin\|^$\|^=\+\|^arch\|^code version\|^host\|^compile_size\|^identifier" > $OUT/sass_$1.txt
  printf "%-34s %-s\n" "sass_$1.txt" "$(c++filt "$fn" | cut -c1-110)" >> $OUT/SUMMARY.txt
  printf "    instructions %s" "$(grep -c '^        /\*[0-9a-f]\{4\}\*/' $OUT/sass_$1.txt)" >> $OUT/SUMMARY.txt
  for op in DFMA UBLKCP UTMALDG LDGMC "STG.*MC\|ST.*\.MC" SYNCS "LDS.128" "STS.128" "RED\." "ATOM" "MEMBAR" "FENCE" USETMAXREG "LDG.E.128" "STG.E.128" "\.STRONG\.SYS"; do
    n=$(grep -c "$op" $OUT/sass_$1.txt || true)
    [ "$n" != "0" ] && printf "  %s=%s" "$op" "$n" >> $OUT/SUMMARY.txt
  done
  echo >> $OUT/SUMMARY.txt
}
: > $OUT/SUMMARY.txt
dump stencil_tstrip_T3          'stencil_tstrip_kernelILi3ELi3E'
dump stencil_tstrip_T2          'stencil_tstrip_kernelILi2ELi2E'
dump stencil_fused_temporal3    'stencil_fused_kernelILi3ELb0ELb0ELi2ELi3ELi64ELi3E'
dump stencil_fused_temporal2    'stencil_fused_kernelILi2ELb0ELb0ELi3ELi2E'
dump stencil_fused_r1_star      'stencil_fused_kernelILi1ELb0ELb0ELi2ELi0E'
dump stencil_strip_r1           'stencil_strip_kernelILi1ELb0ELi5E'
dump halo_push1                 'halo_push1_kernel'
dump land_unpack                'land_unpack_kernel'
dump flag_wait                  'flag_wait_kernel'
dump dot_fused_f64              'dot_fused_kernelIdE'
dump p2p_put_tma                'put_tma_kernel'
dump p2p_put_ldst               'put_ldst_kernel'
dump pingpong_device            'pingpong'
dump pack_move_uniform_8B       'uniform_move_kernelI4int2E'
dump pack_move_uniform_16B      'uniform_move_kernelI4int4E'
dump segment_move_cta           'segment_move_cta_kernel'
dump segment_move_thread        'segment_move_thread_kernel'
dump eager_put                  'eager_put_kernel'
dump eager_copyout              'eager_copyout_kernel'
dump allreduce_oneshot_f32      'allreduce_oneshot_kernelIfE'
dump allreduce_twoshot_f64      'allreduce_twoshot_kernelIdE'
dump allreduce_nvls_f32         'allreduce_nvls_kernelIfE'
dump allreduce_nvls_inplace_f32 'allreduce_nvls_inplace_kernelIfE'
dump coll_push                  'coll11push_kernel'
dump coll_push_multicast        'push_mc_kernel'
dump coll_spread_inplace        'spread_inplace_kernel'
dump coll_pull_local            'pull_local_kernel'
dump coll_barrier               'coll14barrier_kernel'
dump coll_tail                  'coll11tail_kernel'
cat $OUT/SUMMARY.txt
