#!/bin/bash
# N-GPU call: in-kernel collectives / trainers / resume over real peers, then bench.py at N.
N=${1:-2}
mkdir -p gpurun_out
P=$((20000 + RANDOM % 20000))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P tests/multi_worker.py > gpurun_out/multi_worker_n$N.log 2>&1
echo "== multi_worker N=$N exit $?"; grep -E "MULTI OK|Error|error|assert" gpurun_out/multi_worker_n$N.log | head -5; tail -n 3 gpurun_out/multi_worker_n$N.log | cut -c1-300
P=$((20000 + RANDOM % 20000))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --steps 20 --warmup 5 ${BENCH_ARGS:-} > gpurun_out/bench_n$N.log 2>&1
echo "== bench N=$N exit $?"; grep '^{' gpurun_out/bench_n$N.log | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('ours', d['ms_per_step'], d['value'], 'e2e', d['e2e']['value'])
for k in ('nccl_same_compute', 'nccl_baseline'):
    if k in d: print(k, d[k])
print('sim', json.dumps(d.get('sim'))[:600])
" || tail -5 gpurun_out/bench_n$N.log | cut -c1-300
