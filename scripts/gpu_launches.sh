#!/bin/bash
# 1-GPU call: per-launch device time of ONE eager ResNet-50 step (batch $TB200_PROFILE_BATCH) under ncu.
mkdir -p gpurun_out
NCU="ncu --clock-control none --profile-from-start off"
for spec in ${SPECS:-1}; do
  export TB200_TC_CONV=$spec
  timeout 300 python bench/profile_one.py --kernel step > gpurun_out/plain_step_$spec.log 2>&1 && \
    timeout 900 $NCU --metrics gpu__time_duration.sum --csv --log-file gpurun_out/launches_step_$spec.csv \
      python bench/profile_one.py --kernel step > gpurun_out/ncu_step_$spec.log 2>&1
  echo "== launch list TB200_TC_CONV=$spec exit $?"
  python scripts/launch_list.py gpurun_out/launches_step_$spec.csv 45 > gpurun_out/step_launches_$spec.md; head -50 gpurun_out/step_launches_$spec.md
done
