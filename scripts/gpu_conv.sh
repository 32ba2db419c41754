#!/bin/bash
# One 1-GPU gpurun call for the implicit-GEMM convolutions: end-to-end ResNet-50 step with them switched
# on, then an `ncu --set full` capture of one forward convolution (after its plain run exited 0).
mkdir -p gpurun_out
TB200_TC_CONV=1 timeout 400 python bench.py --steps 20 --warmup 5 --skip-nccl > gpurun_out/bench_tcconv.log 2>&1
echo "== bench exit $?"; tail -1 gpurun_out/bench_tcconv.log | cut -c1-400
NCU="ncu --clock-control none --profile-from-start off"
k=conv
timeout 300 python bench/profile_one.py --kernel $k > gpurun_out/plain_$k.log 2>&1 && \
  timeout 600 $NCU --set full --import-source on -f -o gpurun_out/prof_$k \
    python bench/profile_one.py --kernel $k > gpurun_out/ncu_$k.log 2>&1
echo "== ncu $k exit $?"
if [ -f gpurun_out/prof_$k.ncu-rep ]; then
  ncu -i gpurun_out/prof_$k.ncu-rep --page raw --csv > gpurun_out/prof_$k.raw.csv 2>/dev/null
  ncu -i gpurun_out/prof_$k.ncu-rep --page source --csv --print-source sass 2>/dev/null | \
    python scripts/top_sass.py 12 > gpurun_out/prof_$k.top_sass.txt
  [ $(stat -c %s gpurun_out/prof_$k.ncu-rep) -gt 6000000 ] && rm -f gpurun_out/prof_$k.ncu-rep
fi
ls -la gpurun_out | head -30
