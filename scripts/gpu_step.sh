#!/bin/bash
# 1-GPU call: the GPU suites touched by the convolution / BatchNorm work, then the ResNet-50 step with our own
# convolutions (default) and with cuDNN's (TB200_TC_CONV=0), same box, back to back.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_batchnorm.py tests/test_gpu_kernels.py tests/test_gpu_trainer.py -m gpu -q -x > gpurun_out/step_tests.log 2>&1
echo "== tests exit $?"; tail -n 12 gpurun_out/step_tests.log | cut -c1-300
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 3
for spec in ${SPECS:-1 0}; do
  TB200_TC_CONV=$spec timeout 300 python bench.py --steps 20 --warmup 5 --skip-nccl --skip-sim > gpurun_out/bench_conv_$spec.log 2>&1
  echo "== bench TB200_TC_CONV=$spec exit $?"; tail -n 1 gpurun_out/bench_conv_$spec.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read())
    print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], 'loss', d['loss_last'])
except Exception as e:
    print('no json', e)
"
  tail -n 3 gpurun_out/bench_conv_$spec.log | cut -c1-300 | head -2
done
