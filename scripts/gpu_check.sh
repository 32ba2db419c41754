#!/bin/bash
# One 1-GPU gpurun call: every GPU test file (each in its own process, under a timeout), smoke, short bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.csv 2>&1
python -m tiresias_b200.build > gpurun_out/build.log 2>&1 || { echo BUILD FAILED; tail -20 gpurun_out/build.log; exit 1; }
for f in ${TESTS:-tests/test_gpu_*.py tests/test_zz_no_alloc.py}; do
  n=$(basename $f .py)
  timeout 600 python -m pytest $f -m gpu -q -x --timeout 300 > gpurun_out/$n.log 2>&1
  echo "== $n exit $?"; tail -n ${TAIL:-6} gpurun_out/$n.log
done
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "== smoke exit $?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps ${STEPS:-20} --warmup 5 > gpurun_out/bench_n1.log 2>&1; echo "== bench exit $?"; tail -2 gpurun_out/bench_n1.log
