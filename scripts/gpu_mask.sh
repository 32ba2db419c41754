#!/bin/bash
# 1-GPU call: BatchNorm suite (ReLU bit mask), the step with and without the mask, bn_direct, launch list.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_batchnorm.py tests/test_gpu_trainer.py -m gpu -q -x > gpurun_out/mask_tests.log 2>&1
echo "== tests exit $?"; tail -n 15 gpurun_out/mask_tests.log | cut -c1-300
for m in 1 0; do
  TB200_BN_MASK=$m timeout 300 python bench.py --steps 20 --warmup 5 --skip-nccl --skip-sim > gpurun_out/bench_mask$m.log 2>&1
  echo "== bench mask=$m exit $?"; tail -n 1 gpurun_out/bench_mask$m.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read())
    print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], 'loss', d['loss_last'])
except Exception as e:
    print('no json', e)
"
done
timeout 300 python bench/bn_direct.py > gpurun_out/bn_direct_mask.log 2>&1; echo "== bn_direct exit $?"; tail -n 30 gpurun_out/bn_direct_mask.log | cut -c1-250
