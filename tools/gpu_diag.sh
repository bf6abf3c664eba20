#!/bin/bash
# Runs the GPU parity suites group by group, each in its own process (a trapped kernel poisons only its group),
# and leaves one log per group under gpurun_out/.  Usage on the GPU box: bash tools/gpu_diag.sh [groups...]
mkdir -p gpurun_out
: > gpurun_out/diag_summary.txt
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv >> gpurun_out/diag_summary.txt 2>&1
groups=("$@")
if [ ${#groups[@]} -eq 0 ]; then
  groups=(elementwise pack simt tc_gemm tc_conv tc_epilogues attn_simt attn_tc unet)
fi
for grp in "${groups[@]}"; do
  if [ "$grp" = unet ]; then
    timeout 1500 python -m pytest tests/test_unet_gpu.py -m gpu -q -rA --timeout 600 -p no:cacheprovider \
      > gpurun_out/diag_unet.log 2>&1
  else
    timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -rA --timeout 300 -p no:cacheprovider -k "$grp" \
      > gpurun_out/diag_$grp.log 2>&1
  fi
  echo "$grp exit $? : $(tail -n 1 gpurun_out/diag_$grp.log)" >> gpurun_out/diag_summary.txt
done
cat gpurun_out/diag_summary.txt
