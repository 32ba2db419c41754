#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "pull_gemm or gemm" > gpurun_out/pull_test.log 2>&1
echo "== kernel tests exit $?"; tail -n 5 gpurun_out/pull_test.log | cut -c1-300
bash scripts/gpu_n2.sh $N pull multi
