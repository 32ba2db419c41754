#!/bin/bash
mkdir -p gpurun_out
PT="python -m pytest -m gpu -q -x --timeout 200 --timeout-method thread"
timeout 300 $PT tests/test_gpu_norm_pool.py tests/test_gpu_kernels.py -k "pool" > gpurun_out/pool_tests.log 2>&1
echo "== pool tests exit $?"; tail -n 4 gpurun_out/pool_tests.log | cut -c1-300
timeout 300 $PT tests/test_gpu_trainer.py > gpurun_out/pool_tests2.log 2>&1
echo "== trainer tests exit $?"; tail -n 3 gpurun_out/pool_tests2.log | cut -c1-300
timeout 200 python bench.py --steps 20 --warmup 5 --skip-nccl --skip-sim > gpurun_out/bench_pool.log 2>&1
echo "== bench exit $?"; tail -n 1 gpurun_out/bench_pool.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read()); print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'loss', d['loss_last'], d['clocks'])
except Exception as e:
    print('no json', e)
"
TB200_PROFILE_BATCH=256 SPECS="1" bash scripts/gpu_launches.sh | grep -E "launches,|maxpool|stem_pack"
