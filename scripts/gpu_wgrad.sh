#!/bin/bash
# 1-GPU call: halo-tile 3x3 wgrad numerics (bounded), then the convolution table.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_conv.py -m gpu -q -x -k "wgrad_halo or addend or skip_gradient" > gpurun_out/wgrad_test.log 2>&1
rc=$?; echo "== wgrad halo / addend tests exit $rc"; tail -n 15 gpurun_out/wgrad_test.log | cut -c1-300
timeout 600 python bench/kernels.py --only conv --iters 10 > gpurun_out/kernels_conv_wgrad.log 2>&1
echo "== bench exit $?"
python - <<'PY'
import json
for line in open("gpurun_out/kernels_conv_wgrad.log"):
    try:
        d = json.loads(line)
    except Exception:
        continue
    if "shape" in d:
        print(d["note"][:26], {k: v for k, v in d.items() if k.startswith("wgrad") or k in ("fprop_c3_ms", "fprop_cudnn_ms", "dgrad_c3_ms", "dgrad_cudnn_ms")})
PY
