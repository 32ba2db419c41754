#!/usr/bin/env python
"""Kernel microbenchmarks on one B200: every hand-written kernel against its roofline (measured
peaks from MEASURED_PEAKS.json) and against the library op it replaces (cuBLAS via torch.matmul,
torch elementwise SGD). CUDA-event timing, >= 5 warm-ups, L2 flushed (256 MiB write) between timed
iterations. Prints one JSON object per kernel and writes them to gpurun_out/kernels.json.

    python bench/kernels.py [--only gemm|conv|bn|sgd|push|misc] [--iters 20]
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tiresias_b200.ops import native  # noqa: E402
from tiresias_b200.parallel.symm import SymmetricWorld  # noqa: E402


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            d = json.load(fh)
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    except Exception:
        return 6650.0, 1590.0, "fallback"


HBM, TF, SRC = peaks()
_flush = None


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    _flush.fill_(1)


def timeit(fn, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush_l2()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def gemm_cases(iters):
    out = []
    shapes = [  # (M, N, K, note)
        (64, 4096, 25088, "vgg fc6 fwd, batch 64"), (64, 4096, 4096, "vgg fc7 fwd"), (64, 1000, 4096, "vgg fc8 fwd"),
        (256, 4096, 25088, "vgg fc6 fwd, batch 256"), (4096, 25088, 64, "vgg fc6 dW (K = batch)"),
        (64, 1000, 2048, "resnet50 fc fwd"), (8192, 2048, 512, "transformer ffn1"), (8192, 512, 2048, "transformer ffn2"),
        (8192, 1536, 512, "transformer qkv"), (8192, 32000, 512, "transformer lm head"),
        (4096, 4096, 4096, "square 4k"), (8192, 8192, 8192, "square 8k"),
        (802816, 256, 64, "resnet conv3 1x1, layer1 (b256)"), (802816, 64, 256, "resnet conv1 1x1, layer1 (b256)"),
        (200704, 512, 128, "resnet conv3 1x1, layer2 (b256)"), (50176, 1024, 256, "resnet conv3 1x1, layer3 (b256)"),
    ]
    for m, n, k, note in shapes:
        a = torch.randn(m, k, device="cuda").to(torch.bfloat16)
        b = (torch.randn(n, k, device="cuda") / k ** 0.5).to(torch.bfloat16)
        bias = torch.randn(n, device="cuda")
        o = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        ours, ours_min = timeit(lambda: native.gemm_nt(a, b, bias, native.ACT_RELU, out=o), iters)
        ref = torch.relu(a.float() @ b.float().t() + bias)
        err = (o.float() - ref).abs().max().item()
        del ref
        # the same call pinned to the one-CTA kernel and to the 2-CTA kernel (default picks by size)
        native.lib().tb200_gemm_set_pair_mode(0)
        one, _ = timeit(lambda: native.gemm_nt(a, b, bias, native.ACT_RELU, out=o), iters)
        native.lib().tb200_gemm_set_pair_mode(1)
        two, _ = timeit(lambda: native.gemm_nt(a, b, bias, native.ACT_RELU, out=o), iters)
        native.lib().tb200_gemm_set_pair_mode(-1)
        bt = b.t()
        lib, lib_min = timeit(lambda: torch.relu_(torch.addmm(bias.to(torch.bfloat16), a, bt)), iters)
        flops = 2.0 * m * n * k
        byts = 2.0 * (m * k + n * k + m * n)
        roof_ms = max(flops / (TF * 1e12), byts / (HBM * 1e9)) * 1e3
        out.append({"kernel": "gemm_bf16_nt+bias+relu", "shape": [m, n, k], "note": note, "ms": round(ours, 4),
                    "ms_min": round(ours_min, 4), "tflops": round(flops / ours / 1e9, 1),
                    "gbs": round(byts / ours / 1e6, 1), "roofline_ms": round(roof_ms, 4),
                    "frac_of_roofline": round(roof_ms / ours, 3), "cublas_ms": round(lib, 4),
                    "speedup_vs_cublas": round(lib / ours, 3), "one_cta_ms": round(one, 4), "two_cta_ms": round(two, 4),
                    "max_abs_err": round(err, 4), "peaks": SRC})
        print(json.dumps(out[-1]), flush=True)
    return out


def sgd_cases(iters):
    out = []
    for n in (1 << 20, 25 << 20, 138 << 20):
        master = torch.randn(n, device="cuda")
        mom = torch.zeros(n, device="cuda")
        grad = torch.randn(n, device="cuda").to(torch.bfloat16)
        param = torch.empty(n, device="cuda", dtype=torch.bfloat16)
        hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0], device="cuda")
        ours, _ = timeit(lambda: native.sgd_momentum(master, mom, grad, param, hyper), iters)

        def torch_sgd():
            g = grad.float().add_(master, alpha=1e-4)
            mom.mul_(0.9).add_(g)
            master.add_(mom, alpha=-0.1)
            param.copy_(master)
        lib, _ = timeit(torch_sgd, iters)
        byts = n * (4 + 4 + 2 + 4 + 4 + 2)      # read master, mom, grad; write master, mom, param
        out.append({"kernel": "sgd_momentum", "n": n, "ms": round(ours, 4), "gbs": round(byts / ours / 1e6, 1),
                    "frac_of_hbm": round(byts / ours / 1e6 / HBM, 3), "torch_ms": round(lib, 4),
                    "speedup_vs_torch": round(lib / ours, 3), "peaks": SRC})
        print(json.dumps(out[-1]), flush=True)
    return out


def push_cases(iters):
    """tb200_ps_push_pull on one GPU with W emulated ranks sharing the device: flags are pre-posted so
    no kernel waits on another; measures the reduce + SGD + broadcast data path of one PS shard."""
    out = []
    lib = native.lib()
    dev = torch.device("cuda", 0)
    for world, n in ((1, 25 << 20), (8, 25 << 20), (8, 2 << 20)):
        ranks = SymmetricWorld.emulated(world, dev)
        for r in ranks:
            r.alloc("grad", n * 2); r.alloc("param", n * 2); r.alloc("flags", native.FLAG_WORDS * 4)
        for r in ranks:
            r.buffers["grad"].view(torch.bfloat16).normal_()
        shard = n // world
        master = torch.randn(shard, device=dev); mom = torch.zeros(shard, device=dev)
        hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0 / world], device=dev)
        epoch = torch.ones(1, dtype=torch.int32, device=dev)
        counter = torch.zeros(4, dtype=torch.int32, device=dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        me = ranks[0]
        ws = me.native_world("grad", "param", "flags")
        me.post_flags("flags", native.flag_grad_ready(0, 0), native.MAX_WORLD, 1 << 30)

        def run():
            native.check(lib.tb200_ps_push_pull(C.byref(ws), 0, 0, shard, native.BF16, master.data_ptr(),
                                                mom.data_ptr(), hyper.data_ptr(), epoch.data_ptr(),
                                                counter.data_ptr(), status.data_ptr(), 1000, 64, 1,
                                                native.stream_ptr()))
        ms, _ = timeit(run, iters)
        byts = shard * (2 * world + 4 + 4 + 4 + 4 + 2 * world)   # W grad reads, master/mom rw, W param writes
        out.append({"kernel": "ps_push_pull (one shard, emulated ranks on one GPU)", "world": world,
                    "bucket_elems": n, "shard_elems": shard, "ms": round(ms, 4), "gbs": round(byts / ms / 1e6, 1),
                    "frac_of_hbm": round(byts / ms / 1e6 / HBM, 3), "status": int(status.item()), "peaks": SRC})
        print(json.dumps(out[-1]), flush=True)
        for r in ranks:
            r.close()
    return out


def bn_cases(iters):
    """Fused BN(+residual+ReLU) forward and backward on ResNet-50 layer shapes (batch 64) against the
    PyTorch ops they replace (batch_norm + add + relu, autograd backward). Bytes = one pass per tensor the
    op must touch: fwd x (twice: statistics + apply), residual, y; bwd dy, x, y (twice) + dx + dres."""
    from tiresias_b200.ops.batchnorm import fused_bn_act
    out = []
    for n, c, hw, res in ((64, 64, 112, False), (64, 256, 56, True), (64, 128, 28, False), (64, 1024, 14, True),
                          (64, 2048, 7, True), (256, 256, 56, True)):
        shape = (n, c, hw, hw)
        x = torch.randn(shape, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        r = torch.randn_like(x) if res else None
        dy = torch.randn_like(x)
        gamma = torch.ones(c, device="cuda", dtype=torch.bfloat16, requires_grad=True)
        beta = torch.zeros(c, device="cuda", dtype=torch.bfloat16, requires_grad=True)
        rm, rv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
        xg = x.clone().requires_grad_(True)
        rg = r.clone().requires_grad_(True) if res else None
        state = {}

        def fwd():
            state["y"] = fused_bn_act(xg, gamma, beta, rm, rv, rg, True)

        def bwd():
            state["y"].backward(dy, retain_graph=True)
        f_ms, _ = timeit(fwd, iters)
        fwd()
        b_ms, _ = timeit(bwd, iters)
        bn = torch.nn.BatchNorm2d(c).cuda().to(torch.bfloat16).to(memory_format=torch.channels_last)

        def tfwd():
            y = bn(xg)
            if res:
                y = y + rg
            state["ty"] = torch.relu(y)

        def tbwd():
            state["ty"].backward(dy, retain_graph=True)
        tf_ms, _ = timeit(tfwd, iters)
        tfwd()
        tb_ms, _ = timeit(tbwd, iters)
        nbytes = x.numel() * 2
        fbytes = nbytes * (3 + (1 if res else 0))
        bbytes = nbytes * (6 + 1 + (1 if res else 0))
        out.append({"kernel": "fused_bn_relu" + ("_add" if res else ""), "shape": list(shape),
                    "fwd_ms": round(f_ms, 4), "fwd_gbs": round(fbytes / f_ms / 1e6, 1),
                    "fwd_frac_of_hbm": round(fbytes / f_ms / 1e6 / HBM, 3), "bwd_ms": round(b_ms, 4),
                    "bwd_gbs": round(bbytes / b_ms / 1e6, 1), "bwd_frac_of_hbm": round(bbytes / b_ms / 1e6 / HBM, 3),
                    "torch_fwd_ms": round(tf_ms, 4), "torch_bwd_ms": round(tb_ms, 4),
                    "speedup_fwd": round(tf_ms / f_ms, 2), "speedup_bwd": round(tb_ms / b_ms, 2), "peaks": SRC,
                    "note": "L2 flushed between iterations; autograd dispatch overhead included on both sides"})
        print(json.dumps(out[-1]), flush=True)
    return out


def misc_cases(iters):
    out = []
    x = torch.randint(0, 256, (256, 224, 224, 3), dtype=torch.uint8, device="cuda")
    o = torch.empty(x.shape, dtype=torch.bfloat16, device="cuda")
    ms, _ = timeit(lambda: native.u8_to_bf16_norm(x, (0.485, 0.456, 0.406), (0.229, 0.224, 0.225), out=o), iters)
    byts = x.numel() * 3
    out.append({"kernel": "u8_to_bf16_norm", "elems": x.numel(), "ms": round(ms, 4),
                "gbs": round(byts / ms / 1e6, 1), "frac_of_hbm": round(byts / ms / 1e6 / HBM, 3)})
    print(json.dumps(out[-1]), flush=True)
    w = torch.randn(4096, 25088, device="cuda").to(torch.bfloat16)
    ms, _ = timeit(lambda: native.transpose(w), iters)
    byts = w.numel() * 4
    out.append({"kernel": "transpose_bf16", "shape": list(w.shape), "ms": round(ms, 4),
                "gbs": round(byts / ms / 1e6, 1), "frac_of_hbm": round(byts / ms / 1e6 / HBM, 3)})
    print(json.dumps(out[-1]), flush=True)
    return out


def conv_cases(iters):
    """Implicit-GEMM convolutions (im2col TMA + tcgen05) against cuDNN through torch, ResNet-50 shapes at
    batch 256: forward, input gradient (stride 1 only) and weight gradient."""
    out = []
    shapes = [  # (n, cin, cout, hw, k, stride, pad, note)
        (256, 64, 64, 56, 3, 1, 1, "layer1 3x3"), (256, 128, 128, 56, 3, 2, 1, "layer2.0 3x3 s2"),
        (256, 128, 128, 28, 3, 1, 1, "layer2 3x3"), (256, 256, 256, 28, 3, 2, 1, "layer3.0 3x3 s2"),
        (256, 256, 256, 14, 3, 1, 1, "layer3 3x3"), (256, 512, 512, 14, 3, 2, 1, "layer4.0 3x3 s2"),
        (256, 512, 512, 7, 3, 1, 1, "layer4 3x3"), (256, 256, 512, 56, 1, 2, 0, "layer2 downsample 1x1 s2"),
        (256, 512, 1024, 28, 1, 2, 0, "layer3 downsample 1x1 s2"), (256, 1024, 2048, 14, 1, 2, 0, "layer4 downsample 1x1 s2"),
    ]
    aten = torch.ops.aten.convolution_backward
    for n, cin, cout, hw, k, stride, pad, note in shapes:
        x = torch.randn(n, cin, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(cout, cin, k, k, device="cuda") / (cin * k * k) ** 0.5).to(torch.bfloat16)
        w = w.contiguous(memory_format=torch.channels_last)
        y = torch.nn.functional.conv2d(x, w, stride=stride, padding=pad)
        dy = torch.randn_like(y)
        oh = y.shape[2]
        flops = 2.0 * n * oh * oh * cout * cin * k * k
        act_bytes = 2.0 * (x.numel() + y.numel() + w.numel())
        roof_ms = max(flops / (TF * 1e12), act_bytes / (HBM * 1e9)) * 1e3
        row = {"kernel": "conv2d implicit GEMM", "shape": [n, cin, cout, hw, k, stride, pad], "note": note,
               "gflop": round(flops / 1e9, 1), "roofline_ms": round(roof_ms, 4), "peaks": SRC}
        ours, _ = timeit(lambda: native.conv2d_fprop(x, w, stride, pad), iters)
        lib, _ = timeit(lambda: torch.nn.functional.conv2d(x, w, stride=stride, padding=pad), iters)
        err = (native.conv2d_fprop(x, w, stride, pad).float() - y.float()).abs().max().item()
        row.update({"fprop_ms": round(ours, 4), "fprop_cudnn_ms": round(lib, 4), "fprop_frac_of_roofline": round(roof_ms / ours, 3),
                    "fprop_tflops": round(flops / ours / 1e9, 1), "fprop_max_abs_diff_vs_cudnn": round(err, 4)})
        if k == 3 and stride == 1:
            # halo-tile kernel (csrc/conv_halo.cu): input tile fetched once, taps as shifted smem windows
            try:
                halo, _ = timeit(lambda: native.conv2d_fprop_halo(x, w), iters)
                herr = (native.conv2d_fprop_halo(x, w).float() - y.float()).abs().max().item()
                row.update({"fprop_halo_ms": round(halo, 4), "fprop_halo_max_abs_diff_vs_cudnn": round(herr, 4)})
            except Exception as e:  # noqa: BLE001 - bring-up: report, keep benchmarking
                row["fprop_halo_error"] = str(e)[:200]
        if k == 3 and stride == 1 and cout % 64 == 0:
            # second-generation halo-tile kernel (csrc/conv3x3.cu), with and without the BatchNorm-statistics epilogue
            try:
                c3, _ = timeit(lambda: native.conv3x3_fprop(x, w), iters)
                c3s, _ = timeit(lambda: native.conv3x3_fprop(x, w, stats=True), iters)
                c3d, _ = timeit(lambda: native.conv3x3_dgrad(dy, w), iters)
                err3 = (native.conv3x3_fprop(x, w).float() - y.float()).abs().max().item()
                row.update({"fprop_c3_ms": round(c3, 4), "fprop_c3_stats_ms": round(c3s, 4), "dgrad_c3_ms": round(c3d, 4),
                            "fprop_c3_max_abs_diff_vs_cudnn": round(err3, 4), "fprop_c3_tflops": round(flops / c3 / 1e9, 1)})
            except Exception as e:  # noqa: BLE001
                row["fprop_c3_error"] = str(e)[:200]
        if stride == 1:
            ours, _ = timeit(lambda: native.conv2d_dgrad(dy, w, x.shape, stride, pad), iters)
            lib, _ = timeit(lambda: aten(dy, x, w, None, [stride] * 2, [pad] * 2, [1, 1], False, [0, 0], 1,
                                         [True, False, False]), iters)
            row.update({"dgrad_ms": round(ours, 4), "dgrad_cudnn_ms": round(lib, 4)})
        ours, _ = timeit(lambda: native.conv2d_wgrad(x, dy, (k, k), stride, pad), iters)
        lib, _ = timeit(lambda: aten(dy, x, w, None, [stride] * 2, [pad] * 2, [1, 1], False, [0, 0], 1,
                                     [False, True, False]), iters)
        row.update({"wgrad_ms": round(ours, 4), "wgrad_cudnn_ms": round(lib, 4)})
        if k == 3 and stride == 1 and native.conv3x3_wgrad_eligible(x, cout):
            # halo-tile weight gradient (csrc/conv3x3.cu conv3x3_wgrad_kernel)
            try:
                wh, _ = timeit(lambda: native.conv3x3_wgrad(x, dy), iters)
                ref = aten(dy, x, w, None, [1, 1], [1, 1], [1, 1], False, [0, 0], 1, [False, True, False])[1]
                werr = (native.conv3x3_wgrad(x, dy).float() - ref.float()).abs().max().item()
                row.update({"wgrad_halo_ms": round(wh, 4), "wgrad_halo_max_abs_diff_vs_cudnn": round(werr, 4),
                            "wgrad_halo_rel": round(werr / max(ref.float().abs().max().item(), 1e-9), 5)})
            except Exception as e:  # noqa: BLE001
                row["wgrad_halo_error"] = str(e)[:200]
        out.append(row)
        print(json.dumps(row), flush=True)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    native.lib()
    res = []
    for name, fn in (("gemm", gemm_cases), ("conv", conv_cases), ("bn", bn_cases), ("sgd", sgd_cases), ("push", push_cases), ("misc", misc_cases)):
        if args.only and args.only != name:
            continue
        res += fn(args.iters)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "kernels.json"), "w") as fh:
        json.dump(res, fh, indent=1)


if __name__ == "__main__":
    main()
