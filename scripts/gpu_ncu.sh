#!/bin/bash
# One 1-GPU gpurun call: `ncu --set full` captures of the kernels named in $KERNELS (bench/profile_one.py),
# each only after its plain run exited 0. Brings back raw-metric csv + SASS hot spots per kernel.
mkdir -p gpurun_out
NCU="ncu --clock-control none --profile-from-start off"
for k in ${KERNELS:-halo1}; do
  timeout 300 python bench/profile_one.py --kernel $k > gpurun_out/plain_$k.log 2>&1 && \
    timeout 900 $NCU --set full --import-source on -f -o gpurun_out/prof_$k \
      python bench/profile_one.py --kernel $k > gpurun_out/ncu_$k.log 2>&1
  echo "== ncu $k exit $?"
  if [ -f gpurun_out/prof_$k.ncu-rep ]; then
    ncu -i gpurun_out/prof_$k.ncu-rep --page raw --csv > gpurun_out/prof_$k.raw.csv 2>/dev/null
    ncu -i gpurun_out/prof_$k.ncu-rep --page source --csv --print-source sass 2>/dev/null | \
      python scripts/top_sass.py 14 > gpurun_out/prof_$k.top_sass.txt
    [ $(stat -c %s gpurun_out/prof_$k.ncu-rep) -gt 6000000 ] && rm -f gpurun_out/prof_$k.ncu-rep
  fi
done
ls -la gpurun_out | head -40
