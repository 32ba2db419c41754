#!/bin/bash
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 4 --skip-sim 2>&1 | grep '^{' | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('ours', d['ms_per_step'], d['value'], 'fair', d['nccl_same_compute']['ms_per_step'], 'stock', d['nccl_baseline']['ms_per_step'], 'status', d['status'], [ (j['model'], j['ours_ms_per_step']) for j in d.get('communication_bound_jobs', [])])"
