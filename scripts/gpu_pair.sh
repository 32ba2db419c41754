#!/bin/bash
# 1-GPU call: 2-CTA GEMM numerics (bounded by timeouts: a first run of a new barrier protocol), then the GEMM table.
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "pair_gemm_nt_matches_fp32" > gpurun_out/pair_test1.log 2>&1
rc=$?; echo "== pair nt tests exit $rc"; tail -n 12 gpurun_out/pair_test1.log | cut -c1-300
if [ $rc -ne 0 ]; then exit 0; fi
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "gemm" > gpurun_out/pair_test2.log 2>&1
rc=$?; echo "== gemm tests exit $rc"; tail -n 12 gpurun_out/pair_test2.log | cut -c1-300
if [ $rc -ne 0 ]; then exit 0; fi
timeout 600 python bench/kernels.py --only gemm --iters 10 > gpurun_out/kernels_gemm_pair.log 2>&1
echo "== bench exit $?"
python - <<'PY'
import json
for line in open("gpurun_out/kernels_gemm_pair.log"):
    try:
        d = json.loads(line)
    except Exception:
        continue
    if "shape" in d:
        print(d["shape"], d["note"][:28], "default", d["ms"], "one", d.get("one_cta_ms"), "two", d.get("two_cta_ms"), "cublas", d["cublas_ms"], "x", d["speedup_vs_cublas"])
PY
