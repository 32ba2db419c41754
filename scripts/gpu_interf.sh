#!/bin/bash
# 1-GPU: trainer suite (capture mode changed), then co-location through streams (VGG-16 + Transformer in one process).
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_trainer.py -m gpu -q -x --timeout 200 --timeout-method thread > gpurun_out/interf_tests.log 2>&1
echo "== trainer tests exit $?"; tail -n 3 gpurun_out/interf_tests.log | cut -c1-300
timeout 240 python -X faulthandler -m tiresias_b200.calib.interference --models vgg16,transformer --share streams --max-colocated 2 --out gpurun_out/interference_streams.json > gpurun_out/final_interference.log 2>&1
echo "== interference (streams) exit $?"; tail -n 12 gpurun_out/final_interference.log | cut -c1-400
