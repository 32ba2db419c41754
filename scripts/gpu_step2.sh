#!/bin/bash
# 1-GPU call: conv / BN / trainer suites, then the ResNet-50 step (own convolutions) and its launch list.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_batchnorm.py tests/test_gpu_trainer.py -m gpu -q -x > gpurun_out/step_tests.log 2>&1
echo "== tests exit $?"; tail -n 15 gpurun_out/step_tests.log | cut -c1-300
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 3
timeout 300 python bench.py --steps 20 --warmup 5 --skip-nccl --skip-sim > gpurun_out/bench_own.log 2>&1
echo "== bench exit $?"; tail -n 1 gpurun_out/bench_own.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read())
    print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], 'loss', d['loss_last'])
except Exception as e:
    print('no json', e)
"
tail -n 3 gpurun_out/bench_own.log | cut -c1-300 | head -2
TB200_PROFILE_BATCH=256 SPECS="1" bash scripts/gpu_launches.sh
