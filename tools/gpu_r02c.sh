#!/bin/bash
# GPU call r02c: pipelined attention variants + residual-prefetch A/B + full test suite
mkdir -p gpurun_out
python -m pytest tests/test_kernels_gpu.py -x -q -k "attn_tc" 2>&1 | tail -3
python tools/attn_sweep.py np2 pipe2 pipe3 2>&1 | tee gpurun_out/attn_sweep_r02c.txt
LDMB200_LIB=$PWD/dsml_thesis_b200/libldmb200_nopf.so python tools/op_table.py --reps 3 > gpurun_out/op_table_r02c_nopf.txt 2>&1
python tools/op_table.py --reps 3 > gpurun_out/op_table_r02c.txt 2>&1
head -1 gpurun_out/op_table_r02c_nopf.txt gpurun_out/op_table_r02c.txt
grep -h "K160 ks1 s1 res  \|K640 ks1 s1 res f16out\|K320 ks1 s1 res  " gpurun_out/op_table_r02c_nopf.txt | tail -6
echo ---
grep -h "K160 ks1 s1 res  \|K640 ks1 s1 res f16out\|K320 ks1 s1 res  " gpurun_out/op_table_r02c.txt | tail -6
rm -f gpurun_out/parity_report.jsonl
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout 900 > gpurun_out/pytest_gpu_r02c.log 2>&1
echo "pytest exit $? : $(tail -n 1 gpurun_out/pytest_gpu_r02c.log)"
cp gpurun_out/parity_report.jsonl gpurun_out/parity_report_r02c.jsonl 2>/dev/null
