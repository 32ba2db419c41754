#!/usr/bin/env python
"""BatchNorm kernels timed through the C ABI (no autograd, no gradient accumulation): the four passes of
tb200_bn_fwd_train / tb200_bn_bwd on the ResNet-50 layer shapes at batch 256, bytes counted per pass.
TB200_BN_STREAM=0/1 selects the first- / second-generation kernels (read once per process)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tiresias_b200.ops import native  # noqa: E402

HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def timeit(fn, iters=10, warmup=3):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    torch.cuda.set_device(0)
    lib = native.lib()
    dev = torch.device("cuda", 0)
    scratch = torch.zeros(native.BN_SCRATCH_FLOATS, dtype=torch.float32, device=dev)
    st = native.stream_ptr
    for rows, c in ((3211264, 64), (802816, 64), (802816, 256), (200704, 128), (200704, 512), (50176, 256),
                    (50176, 1024), (12544, 512), (12544, 2048)):
        x = torch.randn(rows, c, device=dev).to(torch.bfloat16)
        res = torch.randn(rows, c, device=dev).to(torch.bfloat16)
        y = torch.empty_like(x); dy = torch.randn(rows, c, device=dev).to(torch.bfloat16)
        dx = torch.empty_like(x); dres = torch.empty_like(x)
        gamma = torch.ones(c, device=dev, dtype=torch.bfloat16); beta = torch.zeros(c, device=dev, dtype=torch.bfloat16)
        dg = torch.empty_like(gamma); db = torch.empty_like(gamma)
        rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
        mean, istd = torch.empty(c, device=dev), torch.empty(c, device=dev)

        def fwd(r):
            native.check(lib.tb200_bn_fwd_train(x.data_ptr(), r.data_ptr() if r is not None else 0, y.data_ptr(), gamma.data_ptr(),
                                                beta.data_ptr(), native.BF16, rm.data_ptr(), rv.data_ptr(), 0.1, 1e-5,
                                                mean.data_ptr(), istd.data_ptr(), scratch.data_ptr(), rows, c, 1, st()))

        def bwd(with_y):
            native.check(lib.tb200_bn_bwd(dy.data_ptr(), x.data_ptr(), y.data_ptr() if with_y else 0, gamma.data_ptr(), beta.data_ptr(),
                                          native.BF16, mean.data_ptr(), istd.data_ptr(), dx.data_ptr(),
                                          dres.data_ptr() if with_y else 0, dg.data_ptr(), db.data_ptr(), scratch.data_ptr(),
                                          rows, c, 1, st()))
        mask = torch.zeros(rows * c // 8 + 16, dtype=torch.uint8, device=dev)
        has_mask = lib.tb200_bn_mask_supported(rows, c) == 1

        def fwd_mask():
            native.check(lib.tb200_bn_fwd_train_ex(x.data_ptr(), res.data_ptr(), y.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
                                                   native.BF16, rm.data_ptr(), rv.data_ptr(), 0.1, 1e-5, mean.data_ptr(),
                                                   istd.data_ptr(), scratch.data_ptr(), 0, 0, mask.data_ptr(), rows, c, 1, st()))

        def bwd_mask():
            native.check(lib.tb200_bn_bwd_ex(dy.data_ptr(), x.data_ptr(), 0, mask.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
                                             native.BF16, mean.data_ptr(), istd.data_ptr(), dx.data_ptr(), dres.data_ptr(),
                                             dg.data_ptr(), db.data_ptr(), scratch.data_ptr(), rows, c, 1, st()))
        t = x.numel() * 2
        row = {"rows": rows, "C": c, "mb": round(t / 1e6, 1), "stream": os.environ.get("TB200_BN_STREAM", "1")}
        for name, fn, passes in (("fwd", lambda: fwd(None), 3), ("fwd_res", lambda: fwd(res), 4),
                                 ("bwd_recompute_mask", lambda: bwd(False), 5), ("bwd_res", lambda: bwd(True), 8)):
            ms = timeit(fn)
            row[name + "_ms"] = round(ms, 4)
            row[name + "_frac_hbm"] = round(passes * t / ms / 1e6 / HBM, 3)
        if has_mask:        # residual + ReLU with the mask as bits: forward writes t/16 more, backward reads 2 * t/16 not 2t
            for name, fn, passes in (("fwd_res_bits", fwd_mask, 4 + 1 / 16), ("bwd_res_bits", bwd_mask, 6 + 2 / 16)):
                ms = timeit(fn)
                row[name + "_ms"] = round(ms, 4)
                row[name + "_frac_hbm"] = round(passes * t / ms / 1e6 / HBM, 3)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
