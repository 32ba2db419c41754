#!/bin/bash
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_trainer.py -m gpu -q -x --timeout 100 --timeout-method thread 2>&1 | tail -n 2
timeout 150 python bench.py --steps 5 --warmup 3 --skip-sim 2>&1 | tail -n 1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('ours', d['ms_per_step'], 'fair', d['nccl_same_compute']['ms_per_step'], 'stock', d['nccl_baseline']['ms_per_step'], d['status'])"
