#!/bin/bash
# One 1-GPU gpurun call: kernel microbenchmarks, then (each only after its plain run exited 0) the
# per-launch time list of one training step and `ncu --set full` captures of the hand-written kernels.
mkdir -p gpurun_out
python -m tiresias_b200.build > gpurun_out/build.log 2>&1 || { echo BUILD FAILED; tail -20 gpurun_out/build.log; exit 1; }
timeout 900 python bench/kernels.py --iters 15 > gpurun_out/kernels.log 2>&1; echo "== kernels exit $?"; tail -n 30 gpurun_out/kernels.log
NCU="ncu --clock-control none --profile-from-start off"
timeout 300 python bench/profile_one.py --kernel step > gpurun_out/plain_step.log 2>&1 && \
  timeout 900 $NCU --metrics gpu__time_duration.sum --csv --log-file gpurun_out/launches_step.csv \
    python bench/profile_one.py --kernel step > gpurun_out/ncu_step.log 2>&1
echo "== launch list exit $?"
for k in ${KERNELS:-gemm gemm_splitk push sgd}; do
  timeout 300 python bench/profile_one.py --kernel $k > gpurun_out/plain_$k.log 2>&1 && \
    timeout 900 $NCU --set full --import-source on -f -o gpurun_out/prof_$k \
      python bench/profile_one.py --kernel $k > gpurun_out/ncu_$k.log 2>&1
  echo "== ncu $k exit $?"
  # keep what travels back small (gpurun_out is capped at 64 MiB): raw metrics as csv, SASS-level hot spots
  # as csv, and the .ncu-rep itself only when it is small
  if [ -f gpurun_out/prof_$k.ncu-rep ]; then
    ncu -i gpurun_out/prof_$k.ncu-rep --page raw --csv > gpurun_out/prof_$k.raw.csv 2>/dev/null
    ncu -i gpurun_out/prof_$k.ncu-rep --page source --csv --print-source sass 2>/dev/null | \
      python scripts/top_sass.py 12 > gpurun_out/prof_$k.top_sass.txt
    [ $(stat -c %s gpurun_out/prof_$k.ncu-rep) -gt 6000000 ] && rm -f gpurun_out/prof_$k.ncu-rep
  fi
done
ls -la gpurun_out
