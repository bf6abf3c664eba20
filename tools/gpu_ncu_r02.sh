#!/bin/bash
# ncu --set full of the implicit-GEMM launches of the input half of one eval and of the GroupNorm apply launches.
# Usage on the GPU box: bash tools/gpu_ncu_r02.sh <tag>
tag=${1:-r02}
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
EVAL="python tools/op_table.py --reps 1"
timeout 600 $EVAL > gpurun_out/plain_eval_$tag.log 2>&1 || { echo "plain eval failed"; tail -5 gpurun_out/plain_eval_$tag.log; exit 1; }
cap() {  # cap <name> <kernel regex> <skip> <count>
  timeout 1200 $NCU -k "regex:$2" -s "$3" -c "$4" -o gpurun_out/ncu_$1_$tag $EVAL > gpurun_out/ncu_$1_$tag.log 2>&1
  echo "ncu $1 exit $?"
  if [ -f gpurun_out/ncu_$1_$tag.ncu-rep ]; then
    ncu -i gpurun_out/ncu_$1_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_$1_$tag.csv 2>/dev/null
    [ "$(stat -c %s gpurun_out/ncu_$1_$tag.ncu-rep)" -gt 15000000 ] && rm -f gpurun_out/ncu_$1_$tag.ncu-rep
  fi
}
# three evals run (guarded first call, two timed): skip the first two evals' launches of each family
cap igemm igemm_tc_kernel 294 40
cap gn "gn_apply_kernel" 102 6
ls -la gpurun_out | grep "ncu_.*$tag"
