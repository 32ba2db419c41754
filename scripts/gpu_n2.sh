#!/bin/bash
# N-GPU call: (optionally) the multi-rank checks, the pull-GEMM and push/pull microbenchmarks, bench.py at N.
N=${1:-4}; shift
mkdir -p gpurun_out
port() { echo $((20000 + RANDOM % 20000)); }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for what in "$@"; do
  case $what in
    multi) timeout 900 $TR --master-port $(port) tests/multi_worker.py > gpurun_out/multi_worker_n$N.log 2>&1
           echo "== multi_worker N=$N exit $?"; grep -E "MULTI OK|Error|assert" gpurun_out/multi_worker_n$N.log | head -5; tail -n 2 gpurun_out/multi_worker_n$N.log | cut -c1-300;;
    pull)  timeout 600 $TR --master-port $(port) bench/comm.py --op pull_gemm > gpurun_out/comm_pull_n$N.log 2>&1
           echo "== comm pull_gemm N=$N exit $?"; grep '^{' gpurun_out/comm_pull_n$N.log | cut -c1-400;;
    ps)    timeout 600 $TR --master-port $(port) bench/comm.py --op ps > gpurun_out/comm_ps_n$N.log 2>&1
           echo "== comm ps N=$N exit $?"; grep '^{' gpurun_out/comm_ps_n$N.log | grep -v '"ctas": 148' | cut -c1-330;;
    bench) timeout 900 $TR --master-port $(port) bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2>&1
           echo "== bench N=$N exit $?"; grep '^{' gpurun_out/bench_n$N.log | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('ours', d['ms_per_step'], d['value'], 'e2e', d['e2e']['value'], 'nvls', d.get('nvls'), 'pull', d.get('pull_layers'), 'phases', d.get('phases_ms'))
for k in ('nccl_same_compute', 'nccl_baseline'):
    if k in d: print(k, {a: b for a, b in d[k].items() if a != 'what'})
for r in d.get('communication_bound_jobs', []): print(r)
print('clocks', d.get('clocks'), d.get('nccl_error'))
" || tail -5 gpurun_out/bench_n$N.log | cut -c1-300;;
    benchpull) # the headline with the fc layer pulled inside its GEMM whatever the batch (A/B against `bench`)
           TB200_PULL_MAX_ROWS=1000000 TB200_PULL_MIN_MB=2 timeout 600 $TR --master-port $(port) bench.py --gpus $N --steps 20 --warmup 5 --skip-sim --skip-nccl > gpurun_out/bench_pull_n$N.log 2>&1
           echo "== bench (pull at any batch) N=$N exit $?"; grep '^{' gpurun_out/bench_pull_n$N.log | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('ours+pull', d['ms_per_step'], d['value'], 'e2e', d['e2e']['value'], 'pull', d.get('pull_layers'))
" || tail -5 gpurun_out/bench_pull_n$N.log | cut -c1-300;;
    kernels) timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_multi.py -m gpu -q -x --timeout 240 --timeout-method thread > gpurun_out/kernels_n$N.log 2>&1
           echo "== kernel + multi tests N=$N exit $?"; tail -n 4 gpurun_out/kernels_n$N.log | cut -c1-300;;
  esac
done
