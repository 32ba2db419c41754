#!/bin/bash
# One gpurun --gpus N call: multi-GPU kernel/trainer checks at world N only (NVLS + P2P paths), then the
# communication microbenchmark (P2P vs NVLS vs NCCL) and optionally the N-GPU bench.
N=${1:-2}
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
  --master-port 29541 tests/multi_worker.py > gpurun_out/multi_worker_n$N.log 2>&1
echo "== multi_worker n=$N exit $?"; tail -n 5 gpurun_out/multi_worker_n$N.log | cut -c1-400
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
  --master-port 29542 bench/comm.py > gpurun_out/comm_n$N.log 2>&1
echo "== comm n=$N exit $?"; grep '^{' gpurun_out/comm_n$N.log | cut -c1-330
if [ -n "$BENCH" ]; then
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29543 bench.py --gpus $N --steps 20 --warmup 5 $BENCH > gpurun_out/bench_nvls_n$N.log 2>&1
  echo "== bench n=$N exit $?"; tail -n 2 gpurun_out/bench_nvls_n$N.log | cut -c1-600
fi
