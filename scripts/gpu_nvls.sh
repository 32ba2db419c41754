#!/bin/bash
# One gpurun --gpus N call: NVLS multicast self test (multimem.ld_reduce / multimem.st) at a few sizes.
N=${1:-2}
mkdir -p gpurun_out
timeout 300 python - <<PY > gpurun_out/nvls_n$N.log 2>&1
import json, torch
from tiresias_b200.ops import native
print("devices", torch.cuda.device_count(), "nvls_supported", native.nvls_supported(0))
for mb in (2, 24, 96, 528):
    for ctas in (32, 148):
        try:
            rep = native.nvls_selftest($N, mb << 20, ctas)
        except Exception as e:
            rep = {"exception": str(e)}
        print(json.dumps(rep), flush=True)
        if rep is None:
            print("last_error", native.lib().tb200_last_error().decode())
            break
PY
echo "== nvls exit $?"; cat gpurun_out/nvls_n$N.log | cut -c1-600
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -rA --timeout 500 2>&1 | tail -8
