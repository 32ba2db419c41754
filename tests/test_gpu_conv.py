"""Implicit-GEMM convolutions (tb200_conv2d_fprop / _dgrad / _wgrad: im2col-mode TMA loads feeding the
tcgen05 mainloop) against a plain PyTorch fp32 convolution of the same bf16-rounded operands."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(dev, n, cin, cout, hw, k, stride, pad, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(n, cin, hw, hw, generator=g).to(dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).to(dev).to(torch.bfloat16)
    w = w.contiguous(memory_format=torch.channels_last)
    return x, w


# (n, cin, cout, hw, k, stride, pad): 3x3 same, 3x3 stride 2, 1x1 stride 2, a pixel count that is not a
# multiple of the 128-row tile, an input below the 128 KiB small-tensor threshold, 5x5
SHAPES = [
    (8, 64, 64, 28, 3, 1, 1),
    (8, 128, 128, 28, 3, 2, 1),
    (8, 256, 512, 14, 1, 2, 0),
    (3, 512, 512, 7, 3, 1, 1),
    (2, 64, 128, 14, 3, 1, 1),
    (4, 64, 96, 12, 5, 1, 2),
    (4, 128, 64, 15, 3, 2, 1),
]


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_fprop_matches_fp32(cuda_device, shape):
    from tiresias_b200.ops import native
    n, cin, cout, hw, k, stride, pad = shape
    x, w = _case(cuda_device, *shape)
    y = native.conv2d_fprop(x, w, stride, pad)
    ref = torch.nn.functional.conv2d(x.float(), w.float(), stride=stride, padding=pad)
    assert y.shape == ref.shape and y.is_contiguous(memory_format=torch.channels_last)
    assert torch.allclose(y.float(), ref, rtol=2e-2, atol=2e-2), float((y.float() - ref).abs().max())


# strided input gradients run as stride x stride parity classes (csrc/gemm_tcgen05.cu conv2d_dgrad_strided):
# 3x3 / 2 on an odd image, 7x7 / 2, 1x1 / 2 (three of the four classes are all-zero), 5x5 / 3
STRIDED_DGRAD = [(3, 64, 64, 17, 3, 2, 1), (2, 8, 64, 20, 7, 2, 3), (2, 64, 128, 13, 5, 3, 2), (2, 128, 64, 9, 1, 2, 0)]


@pytest.mark.parametrize("shape", [s for s in SHAPES + STRIDED_DGRAD if s[2] % 64 == 0], ids=lambda s: "x".join(map(str, s)))
def test_dgrad_matches_fp32(cuda_device, shape):
    from tiresias_b200.ops import native
    n, cin, cout, hw, k, stride, pad = shape
    x, w = _case(cuda_device, *shape)
    xr = x.float().requires_grad_(True)
    ref_y = torch.nn.functional.conv2d(xr, w.float(), stride=stride, padding=pad)
    dy = torch.randn_like(ref_y).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ref_y.backward(dy.float())
    dx = native.conv2d_dgrad(dy, w, x.shape, stride, pad)
    assert dx.shape == x.shape
    assert torch.allclose(dx.float(), xr.grad, rtol=2e-2, atol=3e-2), float((dx.float() - xr.grad).abs().max())


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_wgrad_matches_fp32(cuda_device, shape):
    from tiresias_b200.ops import native
    n, cin, cout, hw, k, stride, pad = shape
    x, w = _case(cuda_device, *shape)
    wr = w.float().requires_grad_(True)
    ref_y = torch.nn.functional.conv2d(x.float(), wr, stride=stride, padding=pad)
    dy = (torch.randn_like(ref_y) / 8).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ref_y.backward(dy.float())
    dw = native.conv2d_wgrad(x, dy, (k, k), stride, pad, out_dtype=torch.float32)
    assert dw.shape == w.shape
    scale = float(wr.grad.abs().max())
    assert torch.allclose(dw, wr.grad, rtol=2e-3, atol=2e-3 * scale), float((dw - wr.grad).abs().max())


def test_module_forward_backward(cuda_device):
    from tiresias_b200.ops.conv import TcConv2d
    torch.manual_seed(0)
    for stride in (1, 2):
        conv = TcConv2d(64, 128, 3, stride, 1).to(cuda_device).to(torch.bfloat16).to(memory_format=torch.channels_last)
        x = torch.randn(4, 64, 16, 16, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        x.requires_grad_(True)
        y = conv(x)
        dy = torch.randn_like(y)
        y.backward(dy)
        xr = x.detach().float().requires_grad_(True)
        wr = conv.weight.detach().float().requires_grad_(True)
        yr = torch.nn.functional.conv2d(xr, wr, stride=stride, padding=1)
        yr.backward(dy.float())
        assert torch.allclose(y.float(), yr, rtol=3e-2, atol=3e-2)
        assert torch.allclose(x.grad.float(), xr.grad, rtol=3e-2, atol=6e-2)
        assert conv.weight.grad.shape == conv.weight.shape
        assert torch.allclose(conv.weight.grad.float(), wr.grad, rtol=3e-2, atol=3e-2 * float(wr.grad.abs().max()))


# The first-generation halo-tile kernel (csrc/conv_halo.cu; power-of-two padded rows). Validated on B200 in round 2;
# superseded by csrc/conv3x3.cu (below) and kept as its reference point.
@pytest.mark.parametrize("shape", [(4, 64, 64, 56), (4, 128, 128, 28), (4, 256, 256, 14), (3, 64, 96, 12), (2, 512, 512, 7)],
                         ids=lambda s: "x".join(map(str, s)))
def test_halo_fprop_matches_fp32(cuda_device, shape):
    from tiresias_b200.ops import native
    n, cin, cout, hw = shape
    x, w = _case(cuda_device, n, cin, cout, hw, 3, 1, 1)
    y = native.conv2d_fprop_halo(x, w)
    torch.cuda.synchronize()
    ref = torch.nn.functional.conv2d(x.float(), w.float(), padding=1)
    assert y.shape == ref.shape
    assert torch.allclose(y.float(), ref, rtol=2e-2, atol=2e-2), float((y.float() - ref).abs().max())


# ---- second-generation halo-tile kernel (csrc/conv3x3.cu): flat padded space with pitch W+1 / H+1 ----------
C3_SHAPES = [(4, 64, 64, 56), (4, 128, 128, 28), (4, 256, 256, 14), (2, 512, 512, 7), (3, 64, 96, 12),
             (2, 128, 64, 9), (5, 512, 512, 7), (3, 256, 384, 9), (64, 256, 256, 14), (1, 64, 8, 5), (5, 192, 320, 10),
             (1, 64, 64, 1)]


@pytest.mark.parametrize("shape", C3_SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_conv3x3_fprop_matches_fp32(cuda_device, shape):
    from tiresias_b200.ops import native
    n, cin, cout, hw = shape
    x, w = _case(cuda_device, n, cin, cout, hw, 3, 1, 1)
    y, stats = native.conv3x3_fprop(x, w, stats=True)
    torch.cuda.synchronize()
    ref = torch.nn.functional.conv2d(x.float(), w.float(), padding=1)
    assert y.shape == ref.shape
    assert torch.allclose(y.float(), ref, rtol=1e-2, atol=1e-2), float((y.float() - ref).abs().max())
    # the statistics are those of the bf16 outputs the kernel stored (what a following BatchNorm reads)
    yf = y.float().permute(1, 0, 2, 3).reshape(cout, -1).double()
    sums = stats.sums()
    assert stats.count == n * hw * hw
    assert torch.allclose(sums[:, 0], yf.sum(1), rtol=1e-4, atol=1e-2), float((sums[:, 0] - yf.sum(1)).abs().max())
    assert torch.allclose(sums[:, 1], (yf * yf).sum(1), rtol=1e-4, atol=1e-2)
    y2 = native.conv3x3_fprop(x, w)                       # without the statistics: same output, bit for bit
    assert torch.equal(y2, y)


@pytest.mark.parametrize("shape", C3_SHAPES[:-2], ids=lambda s: "x".join(map(str, s)))
def test_conv3x3_dgrad_matches_fp32(cuda_device, shape):
    from tiresias_b200.ops import native
    n, cin, cout, hw = shape
    if cout % 64:
        pytest.skip("dgrad reduces over cout: needs a multiple of 64")
    x, w = _case(cuda_device, n, cin, cout, hw, 3, 1, 1)
    xr = x.float().requires_grad_(True)
    ref_y = torch.nn.functional.conv2d(xr, w.float(), padding=1)
    dy = (torch.randn_like(ref_y) / 8).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ref_y.backward(dy.float())
    dx = native.conv3x3_dgrad(dy, w)
    torch.cuda.synchronize()
    assert dx.shape == x.shape
    assert torch.allclose(dx.float(), xr.grad, rtol=1e-2, atol=1e-2), float((dx.float() - xr.grad).abs().max())


def test_bn_finalize_from_conv_stats(cuda_device):
    """conv epilogue statistics -> tb200_bn_finalize -> tb200_bn_apply == batch_norm(conv(x)) of torch."""
    import ctypes as C
    from tiresias_b200.ops import native
    n, cin, cout, hw = 8, 64, 128, 14
    x, w = _case(cuda_device, n, cin, cout, hw, 3, 1, 1)
    y, stats = native.conv3x3_fprop(x, w, stats=True)
    gamma = (torch.rand(cout, device=cuda_device) + 0.5).to(torch.bfloat16)
    beta = torch.randn(cout, device=cuda_device).to(torch.bfloat16)
    rm, rv = torch.zeros(cout, device=cuda_device), torch.ones(cout, device=cuda_device)
    mean, invstd = torch.empty(cout, device=cuda_device), torch.empty(cout, device=cuda_device)
    coef = torch.empty(2 * cout, device=cuda_device)
    native.check(native.lib().tb200_bn_finalize(
        stats.partials.data_ptr(), stats.nparts, float(stats.count), cout, gamma.data_ptr(), beta.data_ptr(),
        native.BF16, rm.data_ptr(), rv.data_ptr(), 0.1, 1e-5, mean.data_ptr(), invstd.data_ptr(), coef.data_ptr(),
        native.stream_ptr()))
    torch.cuda.synchronize()
    yf = y.float()
    ref_mean = yf.mean((0, 2, 3))
    ref_var = yf.var((0, 2, 3), unbiased=False)
    assert torch.allclose(mean, ref_mean, rtol=1e-4, atol=1e-5)
    assert torch.allclose(invstd, torch.rsqrt(ref_var + 1e-5), rtol=1e-4, atol=1e-5)
    assert torch.allclose(rm, 0.1 * ref_mean, rtol=1e-4, atol=1e-6)
    assert torch.allclose(rv, 0.9 + 0.1 * yf.var((0, 2, 3), unbiased=True), rtol=1e-4, atol=1e-6)
    assert torch.allclose(coef[:cout], gamma.float() * invstd, rtol=1e-5, atol=1e-6)
    assert torch.allclose(coef[cout:], beta.float() - mean * gamma.float() * invstd, rtol=1e-5, atol=1e-5)


# ---- ResNet stem on the tensor cores (csrc/stem.cu): 7x7 / 2 over 3 channels as a 4x1 conv over a packed tensor ----
@pytest.mark.parametrize("shape", [(2, 32, 32), (3, 37, 45), (2, 224, 224)], ids=lambda s: "x".join(map(str, s)))
def test_packed_stem_matches_fp32_conv(cuda_device, shape):
    from tiresias_b200.ops import native
    from tiresias_b200.ops.conv import packed_stem_conv
    n, h, w = shape
    g = torch.Generator(device="cpu").manual_seed(h * w)
    img = torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8, generator=g).to(cuda_device)
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    weight = (torch.randn(64, 3, 7, 7, generator=g) / 12.0).to(cuda_device).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last).requires_grad_(True)
    packed = native.stem7x7_pack(img, mean, std).permute(0, 3, 1, 2)           # logical [N, 64, OH+3, OW]
    y = packed_stem_conv(packed, weight, emit_stats=True)
    # reference: the same normalised bf16 image through an fp32 convolution
    xn = native.u8_to_bf16_norm(img, mean, std).permute(0, 3, 1, 2).float()
    wr = weight.detach().float().requires_grad_(True)
    ref = torch.nn.functional.conv2d(xn, wr, stride=2, padding=3)
    assert y.shape == ref.shape
    assert torch.allclose(y.float(), ref, rtol=1e-2, atol=2e-2), float((y.float() - ref).abs().max())
    stats = getattr(y, "_tb200_stats", None)
    if stats is not None:                                  # tiny problems are split along K: no statistics
        yf = y.float().permute(1, 0, 2, 3).reshape(64, -1).double()
        assert torch.allclose(stats.sums()[:, 0], yf.sum(1), rtol=1e-4, atol=1e-2)
    dy = (torch.randn_like(ref) / 8).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y.backward(dy)
    ref.backward(dy.float())
    scale = float(wr.grad.abs().max())
    assert weight.grad.shape == weight.shape
    assert torch.allclose(weight.grad.float(), wr.grad, rtol=2e-2, atol=2e-2 * scale), \
        float((weight.grad.float() - wr.grad).abs().max())


@pytest.mark.parametrize("m,n,k", [(802, 256, 64), (3136, 512, 128), (1000, 2048, 512), (256, 264, 72)])
def test_gemm_addend_epilogue(cuda_device, m, n, k):
    """dx = dy . W + skip with the add in the GEMM's epilogue (one- and two-CTA kernels), also in place."""
    from tiresias_b200.ops import native
    torch.manual_seed(m)
    dy = torch.randn(m, k, device=cuda_device).to(torch.bfloat16)
    w = (torch.randn(k, n, device=cuda_device) / k ** 0.5).to(torch.bfloat16)        # [Cout, Cin] as stored
    skip = torch.randn(m, n, device=cuda_device).to(torch.bfloat16)
    ref = dy.float() @ w.float() + skip.float()
    for mode in (0, 1):
        native.lib().tb200_gemm_set_pair_mode(mode)
        try:
            got = native.gemm_add(dy, w, skip, b_mn_major=True)
            assert torch.allclose(got.float(), ref, rtol=1e-2, atol=2e-2), (mode, (got.float() - ref).abs().max().item())
            buf = skip.clone()
            native.gemm_add(dy, w, buf, b_mn_major=True, out=buf)
            assert torch.equal(buf, got)
        finally:
            native.lib().tb200_gemm_set_pair_mode(-1)


@pytest.mark.parametrize("cin,width,stride", [(256, 64, 1), (256, 128, 2), (64, 64, 1)])
def test_bottleneck_skip_gradient_fused_into_conv1_dgrad(cuda_device, cin, width, stride, monkeypatch):
    """The residual block's two input gradients are summed by conv1's dgrad epilogue (no autograd add kernel):
    same gradients as the unfused block."""
    from tiresias_b200.models.vision import Bottleneck
    from tiresias_b200.ops import conv, native
    torch.manual_seed(0)
    blk = Bottleneck(cin, width, stride, fused=True).to(cuda_device).to(torch.bfloat16).to(memory_format=torch.channels_last)
    torch.nn.init.normal_(blk.bn3.weight, 1.0, 0.1)
    x0 = torch.randn(8, cin, 28, 28, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    calls = []
    real = native.gemm_add
    monkeypatch.setattr(native, "gemm_add", lambda *a, **k: (calls.append(1), real(*a, **k))[1])
    outs = []
    for fuse in (True, False):
        monkeypatch.setattr(conv, "FUSE_SKIP_GRAD", fuse)
        blk.zero_grad(set_to_none=True)
        # x is a non-leaf (as inside the network): its gradient is what flows to the previous block
        leaf = x0.clone().requires_grad_(True)
        x = leaf * 1.0
        y = blk(x)
        torch.manual_seed(1)
        y.backward(torch.randn(y.shape, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last))
        outs.append((y.detach().float(), leaf.grad.float(), [p.grad.float() for p in blk.parameters()]))
    assert len(calls) == 1           # the fused pass used the addend epilogue once, the unfused pass never
    (y1, g1, p1), (y0, g0, p0) = outs
    assert torch.equal(y1, y0)
    assert torch.allclose(g1, g0, rtol=2e-2, atol=2e-2 * g0.abs().max().item()), (g1 - g0).abs().max().item()
    for a, b in zip(p1, p0):
        assert torch.allclose(a, b, rtol=2e-2, atol=2e-2 * max(b.abs().max().item(), 1e-3))


@pytest.mark.parametrize("n,hw,cin,cout", [(2, (56, 56), 64, 64), (4, (28, 28), 128, 128), (3, (14, 14), 256, 256),
                                           (5, (7, 7), 512, 512), (2, (17, 13), 64, 128), (1, (9, 30), 128, 64),
                                           (33, (7, 7), 64, 256)])
def test_conv3x3_wgrad_halo_matches_fp32(cuda_device, n, hw, cin, cout):
    """Halo-tile weight gradient (positions as the reduction dimension, taps as shifted MN-major windows)."""
    from tiresias_b200.ops import native
    h, w = hw
    torch.manual_seed(n * 100 + cin)
    x = torch.randn(n, cin, h, w, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dy = (torch.randn(n, cout, h, w, device=cuda_device) / (n * h * w) ** 0.5).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last)
    assert native.conv3x3_wgrad_eligible(x, cout)
    got = native.conv3x3_wgrad(x, dy).float()
    ref = torch.nn.grad.conv2d_weight(x.float(), (cout, cin, 3, 3), dy.float(), stride=1, padding=1)
    assert got.shape == ref.shape
    assert torch.allclose(got, ref, rtol=1e-2, atol=1e-2), (got - ref).abs().max().item()
    again = native.conv3x3_wgrad(x, dy).float()
    assert torch.equal(got, again)                       # fixed split order: bitwise repeatable


WIDE_SHAPES = [(2, 64, 64, 224, 224), (2, 128, 128, 112, 112), (3, 256, 256, 56, 56), (2, 64, 128, 100, 130),
               (1, 64, 64, 9, 300), (2, 192, 384, 13, 13)]


@pytest.mark.parametrize("n,cin,cout,h,w", WIDE_SHAPES)
def test_conv3x3_bias_relu_and_wide_images(cuda_device, n, cin, cout, h, w):
    """Bias + ReLU in the halo kernel's epilogue; images of 100 columns and more run in segment mode (a tile = 128
    columns of one row, stage = three 130-column boxes): forward and input gradient against fp32."""
    from tiresias_b200.ops import native
    torch.manual_seed(h * 1000 + w)
    x = torch.randn(n, cin, h, w, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    wt = (torch.randn(cout, cin, 3, 3, device=cuda_device) / (3.0 * cin ** 0.5)).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last)
    bias = torch.randn(cout, device=cuda_device).to(torch.bfloat16)
    assert native.conv3x3_supported(n, h, w, cin, cout) and native.conv3x3_supported(n, h, w, cout, cin, True)
    ref = torch.nn.functional.conv2d(x.float(), wt.float(), bias.float(), padding=1)
    got = native.conv3x3_fprop_bias(x, wt, bias, True).float()
    assert torch.allclose(got, torch.relu(ref), rtol=2e-2, atol=2e-2), (got - torch.relu(ref)).abs().max().item()
    got = native.conv3x3_fprop_bias(x, wt, None, False).float()
    assert torch.allclose(got, ref - bias.float().view(1, -1, 1, 1), rtol=2e-2, atol=2e-2)
    dy = torch.randn(n, cout, h, w, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dx = native.conv3x3_dgrad(dy, wt).float()
    dref = torch.nn.grad.conv2d_input((n, cin, h, w), wt.float(), dy.float(), padding=1)
    assert torch.allclose(dx, dref, rtol=2e-2, atol=2e-2 * max(1.0, dref.abs().max().item() / 8)), (dx - dref).abs().max().item()


def test_bias_conv_module_matches_torch(cuda_device):
    """TcConv3x3Bias (VGG / AlexNet layers): forward and all three gradients against nn.Conv2d + ReLU in fp32."""
    from tiresias_b200.ops.conv import TcConv3x3Bias
    torch.manual_seed(4)
    mod = TcConv3x3Bias(64, 128, relu=True).to(cuda_device).to(torch.bfloat16).to(memory_format=torch.channels_last)
    x = torch.randn(4, 64, 56, 56, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xa = x.clone().requires_grad_(True)
    y = mod(xa)
    assert type(y.grad_fn).__name__ == "_Conv3x3BiasActFnBackward"
    g = torch.randn_like(y)
    y.backward(g)
    xr = x.float().requires_grad_(True)
    wr = mod.weight.detach().float().requires_grad_(True)
    br = mod.bias.detach().float().requires_grad_(True)
    pre = torch.nn.functional.conv2d(xr, wr, br, padding=1)
    yr = torch.relu(pre).detach()
    # the mask is taken from the bf16 output: push the reference gradient through the same mask
    pre.backward(g.float() * (y.detach() > 0))
    assert torch.allclose(y.float(), yr, rtol=2e-2, atol=2e-2)
    def close(a, b, frac=2e-3):
        bad = (a - b).abs() > 3e-2 * max(1.0, b.abs().max().item() / 4) + 3e-2 * b.abs()
        return bad.float().mean().item() <= frac
    assert close(xa.grad.float(), xr.grad)
    assert close(mod.weight.grad.float(), wr.grad)
    assert close(mod.bias.grad.float(), br.grad)
