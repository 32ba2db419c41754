#!/bin/bash
# One gpurun --gpus N call: one calibration through the worker pool, then the same trace replayed under several
# executed policies, each compared with its simulator counterpart.
#   usage: gpu_replay_multi.sh N JOBS COMM "sched[:extra flags] ..."      e.g. "dlas-gpu sjf gittins fifo:--pack gandiva"
N=${1:-4}; JOBS=${2:-32}; COMM=${3:-ps}; SCHEDS=${4:-"dlas-gpu sjf gittins"}
OUT=gpurun_out/replay2
mkdir -p $OUT
MODELS=${MODELS:-resnet50,vgg16,transformer,alexnet,resnet101}
timeout 600 python -m tiresias_b200.calib.measure --gpus $N --models $MODELS --comm $COMM --iters ${CAL_ITERS:-60} \
  --batch ${BATCH:-32} --out $OUT/calibration_${COMM}_${N}gpu.json --work $OUT/calib_work > $OUT/calib.log 2>&1
echo "== calib exit $?"; tail -n 3 $OUT/calib.log | cut -c1-400
python scripts/make_exec_trace.py $OUT/trace.csv --jobs $JOBS --seed ${SEED:-0} --mean-gap ${GAP:-1.0} --max-iters ${MAX_ITERS:-300}
for entry in $SCHEDS; do
  sched=${entry%%:*}; extra=""
  if [ "$entry" != "$sched" ]; then extra=$(echo "${entry#*:}" | tr ':' ' '); fi
  tag=$(echo "$entry" | tr -c 'A-Za-z0-9\n' '_')
  timeout ${REPLAY_TIMEOUT:-300} python -m tiresias_b200.exec.replay --trace $OUT/trace.csv --gpus $N --schedule $sched --comm $COMM \
    --batch ${BATCH:-32} --calibration $OUT/calibration_${COMM}_${N}gpu.json --out $OUT/run_$tag \
    --queue-limit ${QLIMIT:-20} --check-every ${CHECK_EVERY:-10} $extra > $OUT/replay_$tag.log 2>&1
  echo "== replay $entry exit $?"; tail -n 4 $OUT/replay_$tag.log | cut -c1-500
  rm -rf $OUT/run_$tag/sim
done
rm -rf $OUT/calib_work/calib_run/sim
