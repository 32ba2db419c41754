#!/bin/bash
# One 1-GPU call for everything written while the pod was busy: numerics of every new kernel (each suite bounded,
# so a hang in one does not cost the others), then the step with each new feature switched off in turn, then the
# kernel tables.
mkdir -p gpurun_out
PT="python -m pytest -m gpu -q -x --timeout 240 --timeout-method thread"
run() {  # name, timeout, pytest args...
  local name=$1 to=$2; shift 2
  timeout $to $PT "$@" > gpurun_out/all1_$name.log 2>&1
  local rc=$?
  echo "== $name exit $rc"; tail -n 6 gpurun_out/all1_$name.log | cut -c1-300
  return $rc
}
run bn 600 tests/test_gpu_batchnorm.py
run conv 600 tests/test_gpu_conv.py -k "not wgrad_halo"
run wgrad_halo 300 tests/test_gpu_conv.py -k "wgrad_halo"; wg=$?
run pair_nt 300 tests/test_gpu_kernels.py -k "pair_gemm_nt_matches_fp32"; pr=$?
if [ $pr -eq 0 ]; then run pair_rest 300 tests/test_gpu_kernels.py -k "pair_gemm and not pair_gemm_nt_matches_fp32"; pr=$?; fi
run norm_pool 300 tests/test_gpu_norm_pool.py
run attention 300 tests/test_gpu_attention.py; at=$?
export TB200_OWN_ATTENTION=1
if [ $at -ne 0 ]; then export TB200_OWN_ATTENTION=0; fi
export TB200_HALO_WGRAD=1 TB200_GEMM_2CTA=-1
if [ $wg -ne 0 ]; then export TB200_HALO_WGRAD=0; fi
if [ $pr -ne 0 ]; then export TB200_GEMM_2CTA=0; fi
echo "== features: halo wgrad $TB200_HALO_WGRAD, 2-CTA gemm $TB200_GEMM_2CTA"
run kernels 600 tests/test_gpu_kernels.py -k "not pair_gemm"
run trainer 600 tests/test_gpu_trainer.py
for m in all nomask nofuse nohalo nopair; do
  mask=1; fuse=1; halo=$TB200_HALO_WGRAD; pair=$TB200_GEMM_2CTA
  case $m in nomask) mask=0;; nofuse) fuse=0;; nohalo) halo=0;; nopair) pair=0;; esac
  TB200_BN_MASK=$mask TB200_FUSE_SKIP_GRAD=$fuse TB200_HALO_WGRAD=$halo TB200_GEMM_2CTA=$pair timeout 300 python bench.py --steps 20 --warmup 5 --skip-nccl --skip-sim > gpurun_out/bench_all1_$m.log 2>&1
  echo "== bench $m (mask=$mask fuse=$fuse halo=$halo pair=$pair) exit $?"; tail -n 1 gpurun_out/bench_all1_$m.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read())
    print('ms/step', d['ms_per_step'], 'samples/s', d['value'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], 'loss', d['loss_last'])
except Exception as e:
    print('no json', e)
"
done
timeout 300 python bench/bn_direct.py > gpurun_out/bn_direct_mask.log 2>&1; echo "== bn_direct exit $?"; tail -n 9 gpurun_out/bn_direct_mask.log | cut -c1-500
if [ $pr -eq 0 ]; then
  timeout 600 python bench/kernels.py --only gemm --iters 10 > gpurun_out/kernels_gemm_pair.log 2>&1; echo "== gemm bench exit $?"
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
fi
timeout 600 python bench/kernels.py --only conv --iters 10 > gpurun_out/kernels_conv_wgrad.log 2>&1; echo "== conv bench exit $?"
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
