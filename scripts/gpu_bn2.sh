#!/bin/bash
mkdir -p gpurun_out
for s in 0 1; do
  TB200_BN_STREAM=$s timeout 300 python bench/bn_direct.py > gpurun_out/bn_direct_stream$s.log 2>&1
  echo "== bn direct stream=$s exit $?"; grep '^{' gpurun_out/bn_direct_stream$s.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d['rows'], d['C'], d['mb'], '| fwd', d['fwd_ms'], d['fwd_frac_hbm'], '| fwd_res', d['fwd_res_ms'], d['fwd_res_frac_hbm'], '| bwd2', d['bwd_recompute_mask_ms'], d['bwd_recompute_mask_frac_hbm'], '| bwd_res', d['bwd_res_ms'], d['bwd_res_frac_hbm'])
"
done
