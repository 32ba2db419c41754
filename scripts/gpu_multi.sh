#!/bin/bash
# One gpurun --gpus N call: multi-GPU kernel/trainer checks, then the N-GPU bench (ours + NCCL arm).
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu_n$N.csv 2>&1
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
python -m tiresias_b200.build > gpurun_out/build.log 2>&1 || { echo BUILD FAILED; tail -20 gpurun_out/build.log; exit 1; }
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -rA --timeout 800 > gpurun_out/test_gpu_multi_n$N.log 2>&1
echo "== multi test exit $?"; tail -n 40 gpurun_out/test_gpu_multi_n$N.log
for n in ${BENCH_NS:-1 2 4 8}; do
  [ $n -le $N ] || continue
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_n$n.log 2>&1
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
      --master-port 29511 bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/bench_n$n.log 2>&1
  fi
  echo "== bench n=$n exit $?"; tail -n 3 gpurun_out/bench_n$n.log
done
