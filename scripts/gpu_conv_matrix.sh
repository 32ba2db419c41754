#!/bin/bash
# Which convolution passes are faster on our kernels in situ? One ResNet-50 bench run per routing.
mkdir -p gpurun_out
for spec in "" "p:f" "p:d" "p:w" "k:f" "k:d" "k:w" "1"; do
  TB200_TC_CONV="$spec" timeout 300 python bench.py --steps 15 --warmup 3 --skip-nccl > gpurun_out/bench_conv_tmp.log 2>&1
  echo "spec=[$spec] exit $? $(tail -1 gpurun_out/bench_conv_tmp.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["status"])' 2>&1 | tail -1)"
done | tee gpurun_out/conv_matrix.log
