#!/usr/bin/env python
"""Runs ONE hand-written kernel a few times inside a cudaProfilerStart/Stop range, for
``ncu --profile-from-start off`` (see scripts/gpu_profile.sh). ``--kernel step`` brackets one eager
ResNet-50 PS training step instead (for the per-launch time list)."""
import argparse
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tiresias_b200.ops import native  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kernel", required=True, choices=["gemm", "gemm_splitk", "gemm_smallk", "sgd", "push", "step", "bn", "conv", "halo1", "halo2", "halo3", "c3_1", "c3_2", "c3_3", "c3_4", "wgrad1", "wgrad3", "thin_stats", "stem", "pair8k", "pair_fc6", "wgrad_halo1", "wgrad_halo2", "wgrad_halo3", "attn", "bn_bits", "layernorm"])
    args = ap.parse_args()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    prof = torch.cuda.cudart()
    if args.kernel in ("gemm", "gemm_splitk", "gemm_smallk"):
        m, n, k = {"gemm": (8192, 8192, 8192), "gemm_splitk": (64, 4096, 25088), "gemm_smallk": (8192, 2048, 512)}[args.kernel]
        a = torch.randn(m, k, device=dev).to(torch.bfloat16)
        b = (torch.randn(n, k, device=dev) / k ** 0.5).to(torch.bfloat16)
        bias = torch.randn(n, device=dev)
        o = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
        run = lambda: native.gemm_nt(a, b, bias, native.ACT_RELU, out=o)  # noqa: E731
    elif args.kernel in ("pair8k", "pair_fc6"):
        # the 2-CTA (cta_group::2) kernel, pinned
        m, n, k = {"pair8k": (8192, 8192, 8192), "pair_fc6": (256, 4096, 25088)}[args.kernel]
        native.lib().tb200_gemm_set_pair_mode(1)
        a = torch.randn(m, k, device=dev).to(torch.bfloat16)
        b = (torch.randn(n, k, device=dev) / k ** 0.5).to(torch.bfloat16)
        bias = torch.randn(n, device=dev)
        o = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
        run = lambda: native.gemm_nt(a, b, bias, native.ACT_RELU, out=o)  # noqa: E731
    elif args.kernel.startswith("wgrad_halo"):
        c, hw = {"wgrad_halo1": (64, 56), "wgrad_halo2": (128, 28), "wgrad_halo3": (256, 14)}[args.kernel]
        x = torch.randn(256, c, hw, hw, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        dy = torch.randn(256, c, hw, hw, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        run = lambda: native.conv3x3_wgrad(x, dy)  # noqa: E731
    elif args.kernel == "attn":
        # Transformer-base attention at the bench shape: batch 64 x 128 tokens, 8 heads of 64, causal
        from tiresias_b200.ops.attention import _AttentionFn
        qkv = torch.randn(64 * 128, 1536, device=dev).to(torch.bfloat16).requires_grad_(True)
        dout = torch.randn(64 * 128, 512, device=dev).to(torch.bfloat16)
        run = lambda: _AttentionFn.apply(qkv, 64, 128, 8, True).backward(dout)  # noqa: E731
    elif args.kernel == "layernorm":
        from tiresias_b200.ops.layernorm import FusedLayerNorm
        ln = FusedLayerNorm(512).to(dev).to(torch.bfloat16)
        x = torch.randn(64 * 128 * 16, 512, device=dev).to(torch.bfloat16).requires_grad_(True)
        dy = torch.randn_like(x)
        run = lambda: ln(x).backward(dy)  # noqa: E731
    elif args.kernel == "conv":
        # ResNet-50 layer2 3x3 at batch 256 as an implicit GEMM (im2col-mode TMA loads)
        x = torch.randn(256, 128, 28, 28, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(128, 128, 3, 3, device=dev) / 34.0).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        run = lambda: native.conv2d_fprop(x, w, 1, 1)  # noqa: E731
    elif args.kernel in ("halo1", "halo2", "halo3"):
        # ResNet-50 3x3 / stride 1 convolutions at batch 256 on the halo-tile kernel
        c, hw = {"halo1": (64, 56), "halo2": (128, 28), "halo3": (256, 14)}[args.kernel]
        x = torch.randn(256, c, hw, hw, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(c, c, 3, 3, device=dev) / (3.0 * c ** 0.5)).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        run = lambda: native.conv2d_fprop_halo(x, w)  # noqa: E731
    elif args.kernel.startswith("c3_"):
        # ResNet-50 3x3 / stride 1 convolutions at batch 256 on the second-generation halo-tile kernel
        c, hw = {"c3_1": (64, 56), "c3_2": (128, 28), "c3_3": (256, 14), "c3_4": (512, 7)}[args.kernel]
        x = torch.randn(256, c, hw, hw, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(c, c, 3, 3, device=dev) / (3.0 * c ** 0.5)).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        run = lambda: native.conv3x3_fprop(x, w)  # noqa: E731
    elif args.kernel == "wgrad1":
        # dW of ResNet-50 layer1 conv3 (1x1, 64 -> 256) at batch 256: dy^T . x with both operands as stored
        dy = torch.randn(802816, 256, device=dev).to(torch.bfloat16)
        x = torch.randn(802816, 64, device=dev).to(torch.bfloat16)
        run = lambda: native.gemm(dy, x, a_mn_major=True, b_mn_major=True)  # noqa: E731
    elif args.kernel == "wgrad3":
        x = torch.randn(256, 128, 28, 28, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        dy = torch.randn(256, 128, 28, 28, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        run = lambda: native.conv2d_wgrad(x, dy, (3, 3), 1, 1)  # noqa: E731
    elif args.kernel == "thin_stats":
        # forward of layer1 conv3 with the BatchNorm-statistics epilogue
        x = torch.randn(802816, 64, device=dev).to(torch.bfloat16)
        w = (torch.randn(256, 64, device=dev) / 8).to(torch.bfloat16)
        run = lambda: native.gemm_nt_stats(x, w)  # noqa: E731
    elif args.kernel == "stem":
        img = torch.randint(0, 256, (256, 224, 224, 3), dtype=torch.uint8, device=dev)
        w = (torch.randn(64, 3, 7, 7, device=dev) / 12).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        yp = native.stem7x7_pack(img, (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)).permute(0, 3, 1, 2)
        wp = native.stem7x7_weight_pack(w)
        run = lambda: native.conv2d_fprop_stats(yp, wp, 1, 0)  # noqa: E731
    elif args.kernel in ("bn", "bn_bits"):
        from tiresias_b200.ops.batchnorm import fused_bn_act
        x = torch.randn(64, 256, 56, 56, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
        res = torch.randn_like(x).requires_grad_(True)
        gamma = torch.ones(256, device=dev, dtype=torch.bfloat16, requires_grad=True)
        beta = torch.zeros(256, device=dev, dtype=torch.bfloat16, requires_grad=True)
        rm, rv = torch.zeros(256, device=dev), torch.ones(256, device=dev)
        dy = torch.randn_like(x)
        run = lambda: fused_bn_act(x, gamma, beta, rm, rv, res, True).backward(dy)  # noqa: E731
    elif args.kernel == "sgd":
        n = 25 << 20
        master, mom = torch.randn(n, device=dev), torch.zeros(n, device=dev)
        grad = torch.randn(n, device=dev).to(torch.bfloat16)
        param = torch.empty(n, device=dev, dtype=torch.bfloat16)
        hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0], device=dev)
        run = lambda: native.sgd_momentum(master, mom, grad, param, hyper)  # noqa: E731
    elif args.kernel == "push":
        from tiresias_b200.parallel.symm import SymmetricWorld
        world, n = 8, 25 << 20
        ranks = SymmetricWorld.emulated(world, dev)
        for r in ranks:
            r.alloc("grad", n * 2); r.alloc("param", n * 2); r.alloc("flags", native.FLAG_WORDS * 4)
        shard = n // world
        master, mom = torch.randn(shard, device=dev), torch.zeros(shard, device=dev)
        hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0 / world], device=dev)
        epoch = torch.ones(1, dtype=torch.int32, device=dev)
        counter = torch.zeros(4, dtype=torch.int32, device=dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        ws = ranks[0].native_world("grad", "param", "flags")
        ranks[0].post_flags("flags", native.flag_grad_ready(0, 0), native.MAX_WORLD, 1 << 30)
        lib = native.lib()
        run = lambda: native.check(lib.tb200_ps_push_pull(  # noqa: E731
            C.byref(ws), 0, 0, shard, native.BF16, master.data_ptr(), mom.data_ptr(), hyper.data_ptr(),
            epoch.data_ptr(), counter.data_ptr(), status.data_ptr(), 1000, 64, 1, native.stream_ptr()))
    else:
        from tiresias_b200.models import zoo
        from tiresias_b200.parallel.trainer import Trainer
        spec = zoo.resolve("resnet50")
        bs = int(os.environ.get("TB200_PROFILE_BATCH", "64"))
        tr = Trainer(spec, batch_size=bs, mode="ps", use_graph=False, device=dev)
        x, y = zoo.synthetic_batch(spec, bs)

        def run():
            tr.step(x, y)
            tr.loss()
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    prof.cudaProfilerStart()
    for _ in range(1 if args.kernel == "step" else 3):
        run()
    torch.cuda.synchronize()
    prof.cudaProfilerStop()
    print("profiled", args.kernel)


if __name__ == "__main__":
    main()
