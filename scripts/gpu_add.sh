#!/bin/bash
# 1-GPU: addend epilogue through TMA — numerics, then the step with and without the fused skip gradient.
mkdir -p gpurun_out
PT="python -m pytest -m gpu -q -x --timeout 240 --timeout-method thread"
timeout 300 $PT tests/test_gpu_conv.py -k "addend or skip_gradient" > gpurun_out/add_tests.log 2>&1
echo "== addend tests exit $?"; tail -n 6 gpurun_out/add_tests.log | cut -c1-300
timeout 300 $PT tests/test_gpu_kernels.py -k "gemm" > gpurun_out/add_tests2.log 2>&1
echo "== gemm tests exit $?"; tail -n 4 gpurun_out/add_tests2.log | cut -c1-300
for m in all nofuse; do
  fuse=1; case $m in nofuse) fuse=0;; esac
  TB200_FUSE_SKIP_GRAD=$fuse timeout 300 python bench.py --steps 20 --warmup 5 --skip-nccl --skip-sim > gpurun_out/bench_add_$m.log 2>&1
  echo "== bench $m exit $?"; tail -n 1 gpurun_out/bench_add_$m.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read())
    print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'e2e', d['e2e']['value'], 'loss', d['loss_last'])
except Exception as e:
    print('no json', e)
"
done
TB200_PROFILE_BATCH=256 SPECS="1" bash scripts/gpu_launches.sh | head -24
