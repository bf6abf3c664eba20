#!/bin/bash
# One GPU call: `ncu --set full` captures of every kernel family on the path (sampling and training), exported as
# raw CSV pages for tools/summarize_ncu.py.  Each capture runs only after the same command exited 0 without ncu.
# Usage on the GPU box: bash tools/gpu_ncu_full.sh [tag] [sample|train|both]
tag=${1:-r01}
what=${2:-both}
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"

cap() {  # cap <name> <kernel regex> <skip> <count> <cmd...>
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  timeout 900 $NCU -k "regex:$rx" -s "$skip" -c "$cnt" -o gpurun_out/ncu_${name}_$tag "$@" > gpurun_out/ncu_${name}_$tag.log 2>&1
  local rc=$?
  if [ -f gpurun_out/ncu_${name}_$tag.ncu-rep ]; then
    ncu -i gpurun_out/ncu_${name}_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_${name}_$tag.csv 2>/dev/null
    # gpurun_out/ travels back only below 64 MiB: keep the CSV pages, drop the reports (KEEP_REP=1 keeps small ones)
    if [ -z "$KEEP_REP" ] || [ "$(stat -c %s gpurun_out/ncu_${name}_$tag.ncu-rep)" -gt 12000000 ]; then
      rm -f gpurun_out/ncu_${name}_$tag.ncu-rep
    fi
  fi
  echo "ncu $name exit $rc: $(wc -l < gpurun_out/ncu_${name}_$tag.csv 2>/dev/null) csv lines"
}

if [ "$what" = sample ] || [ "$what" = both ]; then
  EVAL="python tools/op_table.py --reps 1"
  timeout 600 $EVAL > gpurun_out/plain_eval_$tag.log 2>&1 || { echo "plain eval failed"; tail -5 gpurun_out/plain_eval_$tag.log; exit 1; }
  cap igemm   igemm_tc_kernel      148 24 $EVAL    # 24 consecutive launches of the second eval (input half of the UNet)
  cap attn    attention_tc_kernel   16 8  $EVAL    # the 8 attention launches of the input half of the second eval
  cap gn      "gn_apply_kernel|gn_stats_kernel|gn_finalize_kernel" 240 12 $EVAL
  cap ln      layernorm_kernel      72 4  $EVAL
  SHORT="python bench.py --ddim 2 --steps 1 --warmup 1 --graph 0 --no-cpu-baseline"
  timeout 600 $SHORT > gpurun_out/plain_short_$tag.log 2>&1 || { echo "plain short failed"; exit 1; }
  cap ddim    "ddim_step_kernel|small_linear|timestep_embedding|upsample2x|space_to_depth2|nchw_to_nhwc|nhwc_to_nchw" 20 12 $SHORT
fi
if [ "$what" = train ] || [ "$what" = both ]; then
  TRAIN="python bench.py --workload train --hw 32 --steps 1 --warmup 3 --graph 0 --no-cpu-baseline"
  timeout 600 $TRAIN > gpurun_out/plain_train_$tag.log 2>&1 || { echo "plain train failed"; tail -5 gpurun_out/plain_train_$tag.log; exit 1; }
  cap wgrad   wgrad_tc_kernel          170 12 $TRAIN
  cap attnbwd "attention_bwd_tc_kernel|attention_bwd_prep" 48 6 $TRAIN
  cap bwdmem  "gn_bwd_reduce|gn_bwd_apply|layernorm_bwd|grad_cast_kernel|geglu_bwd|adamw_ema" 400 24 $TRAIN
fi
ls -la gpurun_out | grep ncu_ | tail -30
