#!/bin/bash
# One gpurun --gpus N call: calibration of iteration/start-up times through the worker pool, then the
# replay of a mixed trace with simulated-vs-executed JCT (BASELINE.json configs 3-5).
N=${1:-8}; JOBS=${2:-64}; COMM=${3:-ring}; SCHED=${4:-fifo}
mkdir -p gpurun_out/replay
python -m tiresias_b200.build > gpurun_out/build.log 2>&1 || { echo BUILD FAILED; exit 1; }
MODELS=${MODELS:-resnet50,vgg16,transformer,alexnet,resnet101}
timeout 900 python -m tiresias_b200.calib.measure --gpus $N --models $MODELS --comm $COMM --iters ${CAL_ITERS:-80} \
  --batch ${BATCH:-32} --out gpurun_out/replay/calibration_$COMM.json --work gpurun_out/replay/calib_work \
  > gpurun_out/replay/calib.log 2>&1
echo "== calib exit $?"; tail -n 4 gpurun_out/replay/calib.log | cut -c1-600
python scripts/make_exec_trace.py gpurun_out/replay/trace.csv --jobs $JOBS --seed 0 --mean-gap ${GAP:-1.2}
timeout 1200 python -m tiresias_b200.exec.replay --trace gpurun_out/replay/trace.csv --gpus $N --schedule $SCHED --comm $COMM \
  --batch ${BATCH:-32} --calibration gpurun_out/replay/calibration_$COMM.json --out gpurun_out/replay/run_${SCHED}_$COMM \
  --queue-limit ${QLIMIT:-20} --check-every ${CHECK_EVERY:-10} \
  > gpurun_out/replay/replay.log 2>&1
echo "== replay exit $?"; tail -n 6 gpurun_out/replay/replay.log | cut -c1-600
tail -n 6 gpurun_out/replay/run_${SCHED}_$COMM/pool.log | cut -c1-300
rm -rf gpurun_out/replay/calib_work/calib_run/sim gpurun_out/replay/run_${SCHED}_$COMM/sim
