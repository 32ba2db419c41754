#!/bin/bash
# One 1-GPU gpurun call: compute-sanitizer over every hand-written kernel (bench/sanitize_one.py), after the
# plain run has exited 0. Summaries go to gpurun_out/sanitize_<tool>.log (copy into profiles/).
mkdir -p gpurun_out
timeout 300 python bench/sanitize_one.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitize_plain.log; exit 1; }
tail -1 gpurun_out/sanitize_plain.log | cut -c1-300
for tool in ${TOOLS:-memcheck racecheck synccheck}; do
  timeout ${SAN_TIMEOUT:-420} compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 77 \
    python bench/sanitize_one.py > gpurun_out/sanitize_$tool.log 2>&1
  echo "== $tool exit $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|SANITIZE RUN OK|Error|hazard" gpurun_out/sanitize_$tool.log | head -12 | cut -c1-250
done
