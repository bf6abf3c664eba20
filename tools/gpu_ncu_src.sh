#!/bin/bash
# Source-level ncu (stall samples per SASS line) of single implicit-GEMM launches of one eval.
# Usage on the GPU box: bash tools/gpu_ncu_src.sh <tag>
tag=${1:-r02}
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
EVAL="env LDMB200_F16_GUARD=off python tools/op_table.py --reps 1"
# igemm launches of one eval (guard off): conv_in, conv K1440, conv K1440+res, proj_in K160, qkv, to_out+res, geglu, ff.net.2, proj_out+res
# third eval starts after 2 x (147 + 6) igemm launches (two folded upsample convs = 4 launches each)
for spec in ${SPECS:-"conv:307" "projin:309" "toout:311" "geglu:312"}; do
  name=${spec%%:*}; skip=${spec##*:}
  timeout 900 $NCU -k regex:igemm_tc_kernel -s $skip -c 1 -o gpurun_out/ncu_src_${name}_$tag $EVAL > gpurun_out/ncu_src_${name}_$tag.log 2>&1
  echo "ncu $name exit $?"
  ncu -i gpurun_out/ncu_src_${name}_$tag.ncu-rep --page source --csv > gpurun_out/ncu_src_${name}_${tag}_source.csv 2>/dev/null
  ncu -i gpurun_out/ncu_src_${name}_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_src_${name}_${tag}_raw.csv 2>/dev/null
done
ls -la gpurun_out | grep ncu_src
