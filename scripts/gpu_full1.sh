#!/bin/bash
# 1-GPU call: the whole GPU suite, smoke, the full bench line (all arms, simulator comparison, extra models).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/full_tests.log 2>&1
echo "== pytest -m gpu exit $?"; tail -n 6 gpurun_out/full_tests.log | cut -c1-300
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_full_n1.log 2>&1
echo "== bench exit $?"; grep '^{' gpurun_out/bench_full_n1.log | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('ours', d['ms_per_step'], d['value'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], 'phases', d.get('phases_ms'))
for k in ('nccl_same_compute', 'nccl_baseline'):
    if k in d: print(k, {a: b for a, b in d[k].items() if a != 'what'})
for r in d.get('communication_bound_jobs', []): print(r)
print('clocks', d.get('clocks'))
" || tail -5 gpurun_out/bench_full_n1.log | cut -c1-300
timeout 200 python bench.py --impl reference | cut -c1-400
