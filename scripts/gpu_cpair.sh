#!/bin/bash
# 1-GPU: CTA-pair conv3x3 — numerics (bounded), conv table, the step A/B, VGG A/B.
mkdir -p gpurun_out
PT="python -m pytest -m gpu -q -x --timeout 200 --timeout-method thread"
timeout 300 $PT tests/test_gpu_conv.py -k "conv3x3 or bias" > gpurun_out/cpair_tests.log 2>&1
rc=$?; echo "== conv3x3 tests exit $rc"; tail -n 6 gpurun_out/cpair_tests.log | cut -c1-300
if [ $rc -ne 0 ]; then exit 0; fi
for v in 1 0; do
  TB200_CONV_PAIR=$v timeout 200 python bench.py --steps 20 --warmup 5 --skip-nccl --skip-sim > gpurun_out/bench_cpair$v.log 2>&1
  echo "== resnet50 conv_pair=$v exit $?"; tail -n 1 gpurun_out/bench_cpair$v.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read()); print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'loss', d['loss_last'])
except Exception as e:
    print('no json', e)
"
  TB200_CONV_PAIR=$v timeout 200 python bench.py --model vgg16 --batch 128 --steps 10 --warmup 3 --skip-sim --skip-nccl > gpurun_out/bench_vgg_cpair$v.log 2>&1
  echo "== vgg16 conv_pair=$v exit $?"; tail -n 1 gpurun_out/bench_vgg_cpair$v.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read()); print('ms/step', d['ms_per_step'], 'samples/s', d['value'])
except Exception as e:
    print('no json', e)
"
done
timeout 300 python bench/kernels.py --only conv --iters 10 > gpurun_out/kernels_conv_cpair.log 2>&1; echo "== conv bench exit $?"
python - <<'PY'
import json
for line in open("gpurun_out/kernels_conv_cpair.log"):
    try:
        d = json.loads(line)
    except Exception:
        continue
    if "shape" in d and "fprop_c3_ms" in d:
        print(d["note"][:26], {k: v for k, v in d.items() if k in ("fprop_c3_ms", "fprop_c3_stats_ms", "fprop_cudnn_ms", "dgrad_c3_ms", "dgrad_cudnn_ms")})
PY
