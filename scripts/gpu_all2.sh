#!/bin/bash
# 1-GPU follow-up: the suites that failed on the missing import, then the step with each new feature switched off in
# turn, the launch list of the default step and the conv / GEMM tables for the committed profiles.
mkdir -p gpurun_out
PT="python -m pytest -m gpu -q -x --timeout 240 --timeout-method thread"
run() { local name=$1 to=$2; shift 2; timeout $to $PT "$@" > gpurun_out/all2_$name.log 2>&1; local rc=$?; echo "== $name exit $rc"; tail -n 5 gpurun_out/all2_$name.log | cut -c1-300; return $rc; }
run bn 600 tests/test_gpu_batchnorm.py
run conv 600 tests/test_gpu_conv.py
run trainer 600 tests/test_gpu_trainer.py
run kernels 300 tests/test_gpu_kernels.py
for m in all nomask nofuse nohalo nopair allpair; do
  mask=1; fuse=1; halo=1; pair=-1
  case $m in nomask) mask=0;; nofuse) fuse=0;; nohalo) halo=0;; nopair) pair=0;; allpair) pair=1;; esac
  TB200_BN_MASK=$mask TB200_FUSE_SKIP_GRAD=$fuse TB200_HALO_WGRAD=$halo TB200_GEMM_2CTA=$pair timeout 300 python bench.py --steps 20 --warmup 5 --skip-nccl --skip-sim > gpurun_out/bench_all2_$m.log 2>&1
  echo "== bench $m (mask=$mask fuse=$fuse halo=$halo pair=$pair) exit $?"; tail -n 1 gpurun_out/bench_all2_$m.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read())
    print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], 'loss', d['loss_last'])
except Exception as e:
    print('no json', e)
"
done
timeout 200 python bench.py --steps 10 --warmup 3 --model transformer --batch 64 --skip-sim > gpurun_out/bench_all2_transformer.log 2>&1; echo "== transformer bench exit $?"; tail -n 1 gpurun_out/bench_all2_transformer.log | cut -c1-600
TB200_OWN_ATTENTION=0 timeout 200 python bench.py --steps 10 --warmup 3 --model transformer --batch 64 --skip-sim --skip-nccl > gpurun_out/bench_all2_transformer_sdpa.log 2>&1; echo "== transformer (torch SDPA) bench exit $?"; tail -n 1 gpurun_out/bench_all2_transformer_sdpa.log | cut -c1-400
TB200_PROFILE_BATCH=256 SPECS="1" bash scripts/gpu_launches.sh
timeout 300 python bench/kernels.py --only gemm --iters 10 > gpurun_out/kernels_gemm_pair2.log 2>&1; echo "== gemm bench exit $?"
python - <<'PY'
import json
for line in open("gpurun_out/kernels_gemm_pair2.log"):
    try:
        d = json.loads(line)
    except Exception:
        continue
    if "shape" in d:
        print(d["shape"], d["note"][:28], "default", d["ms"], "one", d.get("one_cta_ms"), "two", d.get("two_cta_ms"), "cublas", d["cublas_ms"], "x", d["speedup_vs_cublas"])
PY
