#!/bin/bash
# 1-GPU call: BatchNorm suite (ReLU bit mask), convolution suite (addend epilogue, fused skip gradient, halo wgrad),
# trainer suite, then the ResNet-50 step with each new feature switched off in turn, then bn_direct.
mkdir -p gpurun_out
PT="python -m pytest -m gpu -q -x --timeout 240 --timeout-method thread"
timeout 600 $PT tests/test_gpu_batchnorm.py > gpurun_out/mask_tests_bn.log 2>&1
echo "== bn tests exit $?"; tail -n 6 gpurun_out/mask_tests_bn.log | cut -c1-300
timeout 600 $PT tests/test_gpu_conv.py -k "not wgrad_halo" > gpurun_out/mask_tests_conv.log 2>&1
echo "== conv tests exit $?"; tail -n 8 gpurun_out/mask_tests_conv.log | cut -c1-300
timeout 300 $PT tests/test_gpu_conv.py -k "wgrad_halo" > gpurun_out/mask_tests_wgrad.log 2>&1
wg=$?; echo "== halo wgrad tests exit $wg"; tail -n 8 gpurun_out/mask_tests_wgrad.log | cut -c1-300
export TB200_HALO_WGRAD=1
if [ $wg -ne 0 ]; then export TB200_HALO_WGRAD=0; fi
timeout 600 $PT tests/test_gpu_trainer.py > gpurun_out/mask_tests_trainer.log 2>&1
echo "== trainer tests exit $?"; tail -n 6 gpurun_out/mask_tests_trainer.log | cut -c1-300
for m in all nomask nofuse nohalo; do
  mask=1; fuse=1; halo=$TB200_HALO_WGRAD
  case $m in nomask) mask=0;; nofuse) fuse=0;; nohalo) halo=0;; esac
  TB200_BN_MASK=$mask TB200_FUSE_SKIP_GRAD=$fuse TB200_HALO_WGRAD=$halo timeout 300 python bench.py --steps 20 --warmup 5 --skip-nccl --skip-sim > gpurun_out/bench_mask_$m.log 2>&1
  echo "== bench $m (mask=$mask fuse=$fuse halo=$halo) exit $?"; tail -n 1 gpurun_out/bench_mask_$m.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read())
    print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], 'loss', d['loss_last'])
except Exception as e:
    print('no json', e)
"
done
timeout 300 python bench/bn_direct.py > gpurun_out/bn_direct_mask.log 2>&1; echo "== bn_direct exit $?"; tail -n 12 gpurun_out/bn_direct_mask.log | cut -c1-400
