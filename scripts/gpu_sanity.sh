#!/bin/bash
# Regression check of a finished state: whole GPU suite, smoke(), the bench line with all arms.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 --timeout-method thread > gpurun_out/final_tests.log 2>&1
echo "== pytest -m gpu exit $?"; tail -n 6 gpurun_out/final_tests.log | cut -c1-300
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "== smoke exit $?"; tail -n 2 gpurun_out/final_smoke.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/final_bench.log 2>&1
echo "== bench (no flags) exit $?"; tail -n 1 gpurun_out/final_bench.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('ours', d['ms_per_step'], d['value'], 'steps', d['steps'], 'warmup', d['warmup'], 'e2e', d['e2e'], 'launches', d['gpu_launches'], 'vs_baseline', d['vs_baseline'])
for k in ('nccl_same_compute', 'nccl_baseline'):
    if k in d: print(k, {a: b for a, b in d[k].items() if a != 'what'})
print('sim', [(c['config'], c.get('speedup_wall'), c.get('cluster_csv_identical')) for c in d.get('sim', {}).get('cases', [])])
print('clocks', d.get('clocks'), 'keys', sorted(d.keys()))
" || tail -3 gpurun_out/final_bench.log | cut -c1-300
