#!/bin/bash
# 1-GPU call: numerics of the second-generation 3x3 kernel, then its timing next to cuDNN.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_conv.py -m gpu -q -x -k "conv3x3 or bn_finalize" > gpurun_out/c3_test.log 2>&1
echo "== c3 tests exit $?"; tail -n 12 gpurun_out/c3_test.log | cut -c1-250
TB200_TEST_EXPERIMENTAL=1 timeout 300 python bench/kernels.py --only conv --iters 10 > gpurun_out/kernels_conv_c3.log 2>&1
echo "== conv bench exit $?"; grep '^{' gpurun_out/kernels_conv_c3.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d['note'], '| im2col', d['fprop_ms'], 'halo', d.get('fprop_halo_ms', '-'), 'c3', d.get('fprop_c3_ms', d.get('fprop_c3_error', '-')), 'c3+stats', d.get('fprop_c3_stats_ms', '-'), 'cudnn', d['fprop_cudnn_ms'], '| dgrad c3', d.get('dgrad_c3_ms', '-'), 'im2col', d.get('dgrad_ms', '-'), 'cudnn', d.get('dgrad_cudnn_ms', '-'), '| wgrad', d['wgrad_ms'], 'cudnn', d['wgrad_cudnn_ms'])
"
