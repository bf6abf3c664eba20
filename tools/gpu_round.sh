#!/bin/bash
# One GPU call: full -m gpu suite, the default bench line, a short bench + its ncu launch list.
# Usage on the GPU box: bash tools/gpu_round.sh [tag]
tag=${1:-r01}
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.jsonl
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout 600 > gpurun_out/pytest_gpu_$tag.log 2>&1
echo "pytest exit $? : $(tail -n 1 gpurun_out/pytest_gpu_$tag.log)"
timeout 900 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
echo "bench exit $?"; cat gpurun_out/bench_$tag.json
SHORT="python bench.py --ddim 2 --steps 1 --warmup 1 --graph 0 --no-cpu-baseline"
timeout 600 $SHORT > gpurun_out/bench_short_$tag.json 2> gpurun_out/bench_short_$tag.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 700 -c 700 --csv \
    --log-file gpurun_out/launches_$tag.csv $SHORT > gpurun_out/ncu_$tag.log 2>&1
echo "ncu exit $?"
