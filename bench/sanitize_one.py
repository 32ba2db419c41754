#!/usr/bin/env python
"""Runs every hand-written kernel once on tiny shapes, for compute-sanitizer (scripts/gpu_sanitize.sh):

    compute-sanitizer --tool memcheck|racecheck|synccheck|initcheck python bench/sanitize_one.py

SURVEY.md section 5 (race detection / sanitizers): the reference is single-threaded Python and has
nothing to race; the kernels here (mbarrier pipelines, TMEM, flag protocols) do. Shapes are small because
the tools slow kernels down 10-100x; every code path of the big shapes is still taken (split-K, MN-major
operands, im2col loads, the TMA-store and direct-store epilogues, residual/ReLU BatchNorm variants,
multi-rank push/pull with emulated peers)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tiresias_b200.ops import native  # noqa: E402
from tiresias_b200.ops.batchnorm import fused_bn_act  # noqa: E402
from tiresias_b200.ops.pooling import MaxPool3x3S2  # noqa: E402
from tiresias_b200.parallel.symm import SymmetricWorld  # noqa: E402


def cl(t):
    return t.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)


def main():
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    lib = native.lib()
    done = []

    # GEMM: TMA-store epilogue (+bias+relu), direct-store epilogue (N % 8 != 0), split-K, MN-major operands
    for m, n, k, tag in ((256, 256, 192, "plain"), (96, 100, 128, "direct store"), (64, 256, 4096, "split-K")):
        a = torch.randn(m, k, device=dev).to(torch.bfloat16)
        b = torch.randn(n, k, device=dev).to(torch.bfloat16)
        bias = torch.randn(n, device=dev)
        out = native.gemm_nt(a, b, bias, native.ACT_RELU)
        ref = torch.relu(a.float() @ b.float().t() + bias)
        assert torch.allclose(out.float(), ref, rtol=3e-2, atol=0.5), tag
        done.append("gemm " + tag)
    a = torch.randn(320, 128, device=dev).to(torch.bfloat16)         # [K, M]
    b = torch.randn(320, 192, device=dev).to(torch.bfloat16)         # [K, N]
    out = native.gemm(a, b, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32)
    assert torch.allclose(out, a.float().t() @ b.float(), rtol=3e-2, atol=0.5)
    done.append("gemm MN-major")

    # implicit-GEMM convolutions
    x = cl(torch.randn(2, 64, 10, 10, device=dev))
    w = cl(torch.randn(64, 64, 3, 3, device=dev) / 24)
    y = native.conv2d_fprop(x, w, 1, 1)
    assert torch.allclose(y.float(), torch.nn.functional.conv2d(x.float(), w.float(), padding=1), rtol=3e-2, atol=5e-2)
    dy = cl(torch.randn_like(y))
    native.conv2d_dgrad(dy, w, x.shape, 1, 1)
    native.conv2d_wgrad(x, dy, (3, 3), 1, 1)
    native.conv2d_fprop(x, w, 2, 1)
    done.append("conv fprop/dgrad/wgrad")

    # fused BatchNorm (+residual +ReLU), forward and backward; max pooling
    for residual in (False, True):
        xb = cl(torch.randn(4, 64, 8, 8, device=dev)).requires_grad_(True)
        res = cl(torch.randn(4, 64, 8, 8, device=dev)).requires_grad_(True) if residual else None
        gamma = torch.ones(64, device=dev, dtype=torch.bfloat16, requires_grad=True)
        beta = torch.zeros(64, device=dev, dtype=torch.bfloat16, requires_grad=True)
        rm, rv = torch.zeros(64, device=dev), torch.ones(64, device=dev)
        fused_bn_act(xb, gamma, beta, rm, rv, res, True).backward(cl(torch.randn(4, 64, 8, 8, device=dev)))
    done.append("batchnorm fwd/bwd")
    xp = cl(torch.randn(2, 64, 12, 12, device=dev)).requires_grad_(True)
    MaxPool3x3S2()(xp).sum().backward()
    done.append("maxpool fwd/bwd")

    # optimizer, casts, transposes, norms
    n = 1 << 14
    master, mom = torch.randn(n, device=dev), torch.zeros(n, device=dev)
    grad = torch.randn(n, device=dev).to(torch.bfloat16)
    param = torch.empty(n, device=dev, dtype=torch.bfloat16)
    hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0], device=dev)
    native.sgd_momentum(master, mom, grad, param, hyper)
    native.transpose(torch.randn(96, 160, device=dev).to(torch.bfloat16))
    native.sumsq(grad)
    img = torch.randint(0, 255, (2, 8, 8, 3), dtype=torch.uint8, device=dev)
    native.u8_to_bf16_norm(img, (0.4, 0.4, 0.4), (0.2, 0.2, 0.2))
    done.append("sgd/transpose/sumsq/u8 norm")

    # fused push + SGD + pull and the all-reduce with four emulated ranks on this GPU (flags pre-posted)
    world, n = 4, 1 << 14
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
    native.check(lib.tb200_ps_push_pull(C.byref(ws), 0, 0, shard, native.BF16, master.data_ptr(), mom.data_ptr(),
                                        hyper.data_ptr(), epoch.data_ptr(), counter.data_ptr(), status.data_ptr(),
                                        1_000_000, 8, 1, native.stream_ptr()))
    torch.cuda.synchronize()
    assert int(status.item()) == 0
    done.append("ps_push_pull (4 emulated ranks)")
    one = SymmetricWorld(0, 1, dev)
    one.alloc("buf", n * 2); one.alloc("flags", native.FLAG_WORDS * 4)
    w1 = one.native_world("buf", None, "flags")
    seq = torch.zeros(2, dtype=torch.int32, device=dev)
    ptrs = (C.c_void_p * native.MAX_WORLD)(one.buffers["buf"].ptrs[0])
    for algo in (0, 1):
        native.check(lib.tb200_allreduce(C.byref(w1), ptrs, n, native.BF16, 0.5, algo, seq.data_ptr(),
                                         status.data_ptr(), 1_000_000, 8, native.stream_ptr()))
    torch.cuda.synchronize()
    done.append("allreduce ring/direct (world 1)")
    print("SANITIZE RUN OK:", "; ".join(done), flush=True)


if __name__ == "__main__":
    main()
