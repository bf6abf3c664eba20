#!/bin/bash
# One GPU call: the training-kernel parity tests only (fast iteration while the training step is being built).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_train_kernels_gpu.py -m gpu -q -p no:cacheprovider --timeout 300 -x --no-header -rf "$@" > gpurun_out/pytest_trainkern.log 2>&1
echo "pytest exit $?"; tail -n 40 gpurun_out/pytest_trainkern.log
