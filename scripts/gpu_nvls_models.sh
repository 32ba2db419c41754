#!/bin/bash
# One gpurun --gpus N call: the model benches with the NVLS kernels switched on (TB200_NVLS=1) against the
# peer-load default, same box, back to back — the measurement that decides the default.
N=${1:-8}
mkdir -p gpurun_out
for model in ${MODELS:-resnet50 vgg16}; do
  for nvls in 0 1; do
    TB200_NVLS=$nvls timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port 2955$nvls bench.py --gpus $N --steps 20 --warmup 5 --model $model --skip-nccl \
      > gpurun_out/bench_${model}_n${N}_nvls$nvls.log 2>&1
    echo "== $model n=$N nvls=$nvls exit $? $(tail -n 1 gpurun_out/bench_${model}_n${N}_nvls$nvls.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["value"], d["e2e"]["value"], d["status"])' 2>&1 | tail -n 1)"
  done
done
