#!/bin/bash
# One gpurun --gpus N call: replays of one trace under several executed policies with an EXISTING calibration.
#   usage: gpu_replay_only.sh N CALIBRATION.json "sched[:flag:value...] ..."
N=${1:-4}; CAL=$2; SCHEDS=${3:-"dlas-gpu sjf gittins"}
OUT=gpurun_out/replay3
mkdir -p $OUT
cp $CAL $OUT/calibration.json
python scripts/make_exec_trace.py $OUT/trace.csv --jobs ${JOBS:-16} --seed ${SEED:-1} --mean-gap ${GAP:-1.5} --min-iters ${MIN_ITERS:-300} --max-iters ${MAX_ITERS:-1200}
for entry in $SCHEDS; do
  sched=${entry%%:*}; extra=""
  if [ "$entry" != "$sched" ]; then extra=$(echo "${entry#*:}" | tr ':' ' '); fi
  tag=$(echo "$entry" | tr -c 'A-Za-z0-9\n' '_')
  timeout ${REPLAY_TIMEOUT:-200} python -m tiresias_b200.exec.replay --trace $OUT/trace.csv --gpus $N --schedule $sched --comm ${COMM:-ps} \
    --batch ${BATCH:-32} --calibration $OUT/calibration.json --out $OUT/run_$tag \
    --queue-limit ${QLIMIT:-8} --check-every ${CHECK_EVERY:-10} $extra > $OUT/replay_$tag.log 2>&1
  echo "== replay $entry exit $?"; grep -c "preempt job" $OUT/replay_$tag.log; grep FAILED $OUT/replay_$tag.log | head -2 | cut -c1-200; tail -n 1 $OUT/replay_$tag.log | cut -c1-420
  rm -rf $OUT/run_$tag/sim
done
