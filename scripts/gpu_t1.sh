#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_batchnorm.py tests/test_gpu_kernels.py -m gpu -q > gpurun_out/t1_tests.log 2>&1
echo "== tests exit $?"; tail -n 8 gpurun_out/t1_tests.log | cut -c1-300
KERNELS="wgrad1 thin_stats wgrad3" bash scripts/gpu_ncu.sh 2>&1 | grep "== ncu"
