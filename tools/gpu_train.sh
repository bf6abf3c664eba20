#!/bin/bash
# One GPU call: the training-step tests (kernels + whole step).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_train_kernels_gpu.py tests/test_train_gpu.py -m gpu -q -p no:cacheprovider --timeout 600 --no-header -rf -s "$@" > gpurun_out/pytest_train.log 2>&1
echo "pytest exit $?"; grep -v "^$" gpurun_out/pytest_train.log | tail -n 60
