#!/bin/bash
# One GPU call of round 2: full -m gpu suite, default bench line, op table, ncu --set full of the attention kernel.
# Usage on the GPU box: bash tools/gpu_r02.sh <tag> [pytest -k expression]
tag=${1:-r02}
kexpr=${2:-}
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.jsonl
if [ -n "$kexpr" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout 900 -k "$kexpr" > gpurun_out/pytest_gpu_$tag.log 2>&1
else
  timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout 900 > gpurun_out/pytest_gpu_$tag.log 2>&1
fi
echo "pytest exit $? : $(tail -n 1 gpurun_out/pytest_gpu_$tag.log)"
[ -f gpurun_out/parity_report.jsonl ] && cp gpurun_out/parity_report.jsonl gpurun_out/parity_report_$tag.jsonl
timeout 900 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
echo "bench exit $?"; cut -c1-1500 gpurun_out/bench_$tag.json
timeout 600 python tools/op_table.py --reps 3 > gpurun_out/op_table_$tag.txt 2> gpurun_out/op_table_$tag.err
echo "op_table exit $?"; head -1 gpurun_out/op_table_$tag.txt
timeout 300 python tools/attn_one.py 64 4096 5 3 > gpurun_out/attn_one_$tag.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attention_tc -s 2 -c 1 -f \
    -o gpurun_out/ncu_attn_$tag python tools/attn_one.py 64 4096 5 3 > gpurun_out/ncu_attn_$tag.log 2>&1
echo "ncu attn exit $?"
if [ -f gpurun_out/ncu_attn_$tag.ncu-rep ]; then
  ncu -i gpurun_out/ncu_attn_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_attn_${tag}_raw.csv 2>/dev/null
  ncu -i gpurun_out/ncu_attn_$tag.ncu-rep --page source --csv > gpurun_out/ncu_attn_${tag}_source.csv 2>/dev/null
  ls -la gpurun_out/ncu_attn_$tag.ncu-rep
fi
