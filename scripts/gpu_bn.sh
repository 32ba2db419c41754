#!/bin/bash
# 1-GPU call: numerics of the streamed BatchNorm kernels, then old vs new timing on ResNet-50 layer shapes.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_batchnorm.py -m gpu -q -x > gpurun_out/bn_test.log 2>&1
echo "== bn tests exit $?"; tail -n 15 gpurun_out/bn_test.log | cut -c1-250
for s in 0 1; do
  TB200_BN_STREAM=$s timeout 300 python bench/kernels.py --only bn --iters 15 > gpurun_out/kernels_bn_stream$s.log 2>&1
  echo "== bn bench stream=$s exit $?"
  grep '^{' gpurun_out/kernels_bn_stream$s.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d['kernel'], d['shape'], 'fwd', d['fwd_ms'], d['fwd_frac_of_hbm'], 'bwd', d['bwd_ms'], d['bwd_frac_of_hbm'], 'torch', d['torch_fwd_ms'], d['torch_bwd_ms'])
"
done
