#!/bin/bash
# First GPU call of the next round (1 GPU): bring up what round 1 left unvalidated, each step under its own
# timeout so a hang costs seconds, not the call.
#   1. halo-tile convolution (csrc/conv_halo.cu): numerics against fp32, then timing next to cuDNN and the
#      im2col-TMA kernel
#   2. the whole 1-GPU suite, smoke, bench
mkdir -p gpurun_out
export TB200_TEST_EXPERIMENTAL=1
timeout 120 python -m pytest tests/test_gpu_conv.py tests/test_gpu_kernels.py -m gpu -q -k "halo or pull_gemm" > gpurun_out/halo_test.log 2>&1
echo "== halo tests exit $?"; tail -n 15 gpurun_out/halo_test.log | cut -c1-300
timeout 60 python bench/probe_shift.py > gpurun_out/probe_shift.log 2>&1
echo "== probes exit $?"; grep -E "MN-major|tma_dst" gpurun_out/probe_shift.log | cut -c1-160
timeout 300 python bench/kernels.py --only conv --iters 10 > gpurun_out/kernels_conv_halo.log 2>&1
echo "== conv bench exit $?"; grep '^{' gpurun_out/kernels_conv_halo.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d['note'], 'ours', d['fprop_ms'], 'halo', d.get('fprop_halo_ms', d.get('fprop_halo_error', '-')), 'cudnn', d['fprop_cudnn_ms'])
"
unset TB200_TEST_EXPERIMENTAL
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -n 5
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.log 2>&1; tail -n 1 gpurun_out/bench_n1.log | cut -c1-600
