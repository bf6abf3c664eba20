#!/bin/bash
# One GPU call: op table (CUDA events per launch) + ncu --set full captures of the attention and implicit-GEMM kernels.
# Usage on the GPU box: bash tools/gpu_prof.sh [tag] [attn|gemm|both]
tag=${1:-r01}
what=${2:-both}
mkdir -p gpurun_out
CMD="python tools/op_table.py --reps 1"
timeout 600 $CMD > gpurun_out/op_table_$tag.txt 2> gpurun_out/op_table_$tag.err || { echo "op_table failed"; tail -5 gpurun_out/op_table_$tag.err; exit 1; }
if [ "$what" = attn ] || [ "$what" = both ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attention_tc -s 16 -c 3 \
    -f -o gpurun_out/prof_attn_$tag $CMD > gpurun_out/ncu_attn_$tag.log 2>&1
echo "ncu attn exit $?"
fi
if [ "$what" = gemm ] || [ "$what" = both ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:igemm_tc -s 148 -c 8 \
    -f -o gpurun_out/prof_gemm_$tag $CMD > gpurun_out/ncu_gemm_$tag.log 2>&1
echo "ncu gemm exit $?"
fi
if [ "$what" = traffic ]; then
# DRAM bytes of every implicit-GEMM launch of ONE UNet eval (147 launches; the first eval's 148 are skipped)
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:igemm_tc -s 148 -c 147 --csv --log-file gpurun_out/traffic_gemm_$tag.csv $CMD > gpurun_out/ncu_traffic_$tag.log 2>&1
echo "ncu traffic exit $?"
fi
ls -la gpurun_out | tail -8
