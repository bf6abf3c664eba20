#!/bin/bash
# ONE multi-GPU call (gpurun --gpus N): the headline bench with the strong-scaling key, the C5 sweep (both readings),
# the C3 progressive workload and the C4 training step at N GPUs.  Usage: bash tools/gpu_n8.sh <tag> <N>
tag=${1:-r02}
N=${2:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n${N}_$tag.json 2> gpurun_out/bench_n${N}_$tag.err
echo "bench n$N exit $?"; cut -c1-300 gpurun_out/bench_n${N}_$tag.json
timeout 900 $TR bench.py --gpus $N --workload c5 --steps 1 --warmup 1 > gpurun_out/bench_c5_n${N}_$tag.jsonl 2> gpurun_out/bench_c5_n${N}_$tag.err
echo "c5 exit $?"; wc -l gpurun_out/bench_c5_n${N}_$tag.jsonl
timeout 600 $TR bench.py --gpus $N --workload train --hw 32 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_train32_n${N}_$tag.json 2> gpurun_out/bench_train32_n${N}_$tag.err
echo "train32 exit $?"; cut -c1-200 gpurun_out/bench_train32_n${N}_$tag.json
timeout 600 $TR bench.py --gpus $N --workload train --hw 64 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_train64_n${N}_$tag.json 2> gpurun_out/bench_train64_n${N}_$tag.err
echo "train64 exit $?"; cut -c1-200 gpurun_out/bench_train64_n${N}_$tag.json
timeout 600 $TR bench.py --gpus $N --workload c3 --first-stage 1 --steps 1 --warmup 1 > gpurun_out/bench_c3_n${N}_$tag.json 2> gpurun_out/bench_c3_n${N}_$tag.err
echo "c3 exit $?"; cut -c1-200 gpurun_out/bench_c3_n${N}_$tag.json
# the automatic gradient sync of the drop-in under a multi-rank process group (no explicit GradAllReduce in user code)
timeout 300 $TR tools/check_auto_sync.py > gpurun_out/auto_sync_n${N}_$tag.log 2>&1
echo "auto-sync exit $?"; tail -2 gpurun_out/auto_sync_n${N}_$tag.log
