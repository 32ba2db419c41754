#!/bin/bash
# 1-GPU: statistics epilogue on eight warps — numerics, then the step A/B.
mkdir -p gpurun_out
PT="python -m pytest -m gpu -q -x --timeout 240 --timeout-method thread"
timeout 400 $PT tests/test_gpu_conv.py tests/test_gpu_batchnorm.py > gpurun_out/epi8_tests.log 2>&1
echo "== conv + bn tests exit $?"; tail -n 6 gpurun_out/epi8_tests.log | cut -c1-300
timeout 300 $PT tests/test_gpu_trainer.py tests/test_gpu_kernels.py > gpurun_out/epi8_tests2.log 2>&1
echo "== trainer + kernel tests exit $?"; tail -n 4 gpurun_out/epi8_tests2.log | cut -c1-300
for m in epi8 epi4; do
  v=1; case $m in epi4) v=0;; esac
  TB200_STATS_EPI8=$v timeout 300 python bench.py --steps 20 --warmup 5 --skip-nccl --skip-sim > gpurun_out/bench_$m.log 2>&1
  echo "== bench $m exit $?"; tail -n 1 gpurun_out/bench_$m.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read())
    print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'e2e', d['e2e']['value'], 'loss', d['loss_last'])
except Exception as e:
    print('no json', e)
"
done
TB200_PROFILE_BATCH=256 SPECS="1" bash scripts/gpu_launches.sh | head -16
