#!/bin/bash
# 1-GPU: packing replay (every job normalised to one GPU, up to three tenants share it through streams).
mkdir -p gpurun_out/replay4
python scripts/make_exec_trace.py gpurun_out/replay4/trace.csv --jobs 8 --seed 5 --mean-gap 1.0 --min-iters 300 --max-iters 700
timeout 150 python -m tiresias_b200.exec.replay --trace gpurun_out/replay4/trace.csv --gpus 1 --schedule fifo --comm ps --batch 32 \
  --calibration profiles/replay_r2b/calibration.json --out gpurun_out/replay4/run_pack --pack --max-share 3 --interference 1.0,1.57,2.18 > gpurun_out/replay4/replay_pack.log 2>&1
echo "== pack replay exit $?"; grep FAILED gpurun_out/replay4/replay_pack.log | head -3 | cut -c1-200; tail -n 1 gpurun_out/replay4/replay_pack.log | cut -c1-420
rm -rf gpurun_out/replay4/run_pack/sim
