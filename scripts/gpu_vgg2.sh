#!/bin/bash
mkdir -p gpurun_out
for w in 56 112 28; do
  TB200_BIAS_CONV_MAX_W=$w timeout 240 python bench.py --model vgg16 --batch 128 --steps 10 --warmup 3 --skip-sim --skip-nccl > gpurun_out/bench_vgg16_w$w.log 2>&1
  echo "== vgg16 max_w=$w exit $?"; tail -n 1 gpurun_out/bench_vgg16_w$w.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read()); print('ms/step', d['ms_per_step'], 'samples/s', d['value'])
except Exception as e:
    print('no json', e)
"
done
