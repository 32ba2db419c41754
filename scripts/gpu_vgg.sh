#!/bin/bash
# 1-GPU: VGG-16 / AlexNet steps with the 3x3 layers on our halo kernels vs cuDNN; three tenants through streams.
mkdir -p gpurun_out
for model in vgg16:128 alexnet:256; do
  m=${model%%:*}; b=${model#*:}
  for own in 1 0; do
    TB200_OWN_BIAS_CONV=$own timeout 240 python bench.py --model $m --batch $b --steps 10 --warmup 3 --skip-sim --skip-nccl > gpurun_out/bench_${m}_own$own.log 2>&1
    echo "== $m own_bias_conv=$own exit $?"; tail -n 1 gpurun_out/bench_${m}_own$own.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read()); print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'loss', d.get('loss_last'))
except Exception as e:
    print('no json', e)
"
  done
done
timeout 240 python -m tiresias_b200.calib.interference --models vgg16,transformer --share streams --max-colocated 3 --out gpurun_out/interference_streams3.json > gpurun_out/interference3.log 2>&1
echo "== interference (3 tenants) exit $?"; tail -n 2 gpurun_out/interference3.log | cut -c1-400
