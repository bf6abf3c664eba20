#!/bin/bash
# Final evidence of a round: default bench line, op table, ncu launch list of a short bench, per-launch DRAM traffic of one
# eval (all kernel families), full -m gpu suite with the parity report.  Usage on the GPU box: bash tools/gpu_final.sh <tag>
tag=${1:-r02}
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.jsonl
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout 900 > gpurun_out/pytest_gpu_$tag.log 2>&1
echo "pytest exit $? : $(tail -n 1 gpurun_out/pytest_gpu_$tag.log)"
cp gpurun_out/parity_report.jsonl gpurun_out/parity_report_$tag.jsonl 2>/dev/null
timeout 900 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
echo "bench exit $?"; cut -c1-300 gpurun_out/bench_$tag.json
timeout 600 python tools/op_table.py --reps 3 > gpurun_out/op_table_$tag.txt 2> gpurun_out/op_table_$tag.err
echo "op_table exit $?"; head -1 gpurun_out/op_table_$tag.txt
RX='regex:igemm|attention|gn_|layernorm|small_linear|timestep|nchw|nhwc|space_to_depth|upsample|cast_rows'
LDMB200_F16_GUARD=off timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k "$RX" --csv --log-file gpurun_out/traffic_all_$tag.csv python tools/op_table.py --reps 1 > gpurun_out/ncu_traffic_$tag.log 2>&1
echo "ncu traffic exit $?: $(grep -c '^"' gpurun_out/traffic_all_$tag.csv) csv rows"
SHORT="python bench.py --ddim 2 --steps 1 --warmup 1 --graph 0 --no-cpu-baseline"
timeout 600 $SHORT > gpurun_out/bench_short_$tag.json 2> gpurun_out/bench_short_$tag.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
    --log-file gpurun_out/launches_$tag.csv $SHORT > gpurun_out/ncu_launches_$tag.log 2>&1
echo "ncu launches exit $?"
timeout 300 python bench.py --impl reference --steps 10 --warmup 1 > gpurun_out/bench_reference_$tag.json 2>&1
echo "reference arm exit $?"; cut -c1-300 gpurun_out/bench_reference_$tag.json
