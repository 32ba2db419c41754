"""Fused BatchNorm(+residual)(+ReLU) kernels against a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

from tiresias_b200.ops.batchnorm import FusedBatchNormAct, bn_apply, fused_bn_act

pytestmark = pytest.mark.gpu

SHAPES = [(4, 2048, 7, 7), (2, 64, 56, 56), (3, 1024, 14, 14), (2, 72, 17, 13), (8, 256, 28, 28), (1, 8, 5, 3)]


def _reference(x, res, gamma, beta, relu, eps=1e-5):
    xf = x.float().detach().requires_grad_(True)
    rf = res.float().detach().requires_grad_(True) if res is not None else None
    g = gamma.float().detach().requires_grad_(True)
    b = beta.float().detach().requires_grad_(True)
    rm, rv = torch.zeros_like(g), torch.ones_like(g)
    y = torch.nn.functional.batch_norm(xf, rm, rv, g, b, True, 0.1, eps)
    if rf is not None:
        y = y + rf
    if relu:
        y = torch.relu(y)
    return xf, rf, g, b, y, rm, rv


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("relu,with_res", [(False, False), (True, False), (True, True)])
def test_forward_backward_match_fp32_reference(cuda_device, shape, relu, with_res):
    n, c, h, w = shape
    g0 = torch.Generator(device=cuda_device).manual_seed(n * 1000 + c)
    x = (torch.randn(shape, device=cuda_device, generator=g0) * 1.7 + 0.4).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last).requires_grad_(True)
    res = torch.randn(shape, device=cuda_device, generator=g0).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last).requires_grad_(True) if with_res else None
    gamma = (torch.rand(c, device=cuda_device, generator=g0) + 0.5).to(torch.bfloat16).requires_grad_(True)
    beta = (torch.randn(c, device=cuda_device, generator=g0) * 0.3).to(torch.bfloat16).requires_grad_(True)
    rm = torch.zeros(c, device=cuda_device)
    rv = torch.ones(c, device=cuda_device)
    y = fused_bn_act(x, gamma, beta, rm, rv, res, relu)
    dy = torch.randn(shape, device=cuda_device, generator=g0).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last)
    y.backward(dy)

    xf, rf, gf, bf, yf, rmf, rvf = _reference(x, res, gamma, beta, relu)
    yf.backward(dy.float())
    assert torch.allclose(y.float(), yf, rtol=2e-2, atol=3e-2), (y.float() - yf).abs().max().item()
    assert torch.allclose(rm, rmf, rtol=1e-3, atol=1e-4) and torch.allclose(rv, rvf, rtol=1e-3, atol=1e-4)
    # the ReLU mask is taken from the bf16 output; elements whose fp32 pre-activation is within
    # rounding of zero may flip, so compare gradients with a small outlier allowance
    def close(a, b, rtol, atol, frac=2e-3):
        bad = (a - b).abs() > atol + rtol * b.abs()
        return bad.float().mean().item() <= frac
    assert close(x.grad.float(), xf.grad, 3e-2, 3e-2), (x.grad.float() - xf.grad).abs().max().item()
    if with_res:
        assert close(res.grad.float(), rf.grad, 2e-2, 2e-2)
    rows = n * h * w
    tol = 2e-2 * max(1.0, rows ** 0.5)
    # one flipped ReLU-mask element moves a whole channel sum by O(1): allow a few outlier channels
    frac = 2.0 / c if relu else 0.0
    assert close(gamma.grad.float(), gf.grad, 3e-2, tol, frac), (gamma.grad.float() - gf.grad).abs().max()
    assert close(beta.grad.float(), bf.grad, 3e-2, tol, frac), (beta.grad.float() - bf.grad).abs().max()


def test_bitwise_repeatable_and_scratch_reuse(cuda_device):
    x = torch.randn(8, 256, 28, 28, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    gamma = torch.ones(256, device=cuda_device, dtype=torch.bfloat16)
    beta = torch.zeros(256, device=cuda_device, dtype=torch.bfloat16)
    outs = []
    for _ in range(3):
        rm, rv = torch.zeros(256, device=cuda_device), torch.ones(256, device=cuda_device)
        outs.append((fused_bn_act(x, gamma, beta, rm, rv, None, True).clone(), rm.clone(), rv.clone()))
        # a different layer shape in between shares the same scratch
        small = torch.randn(2, 64, 9, 9, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        fused_bn_act(small, gamma[:64].contiguous(), beta[:64].contiguous(), None, None, None, False)
    for y, rm, rv in outs[1:]:
        assert torch.equal(y, outs[0][0]) and torch.equal(rm, outs[0][1]) and torch.equal(rv, outs[0][2])


@pytest.mark.parametrize("shape", [(32, 256, 28, 28), (16, 64, 56, 56), (64, 2048, 7, 7), (24, 512, 28, 28)])
def test_relu_bit_mask_equals_saved_output_path(cuda_device, shape, monkeypatch):
    """Residual + ReLU layers save one bit per element instead of y: same gradients, bit for bit."""
    from tiresias_b200.ops import batchnorm, native
    n, c, h, w = shape
    assert native.lib().tb200_bn_mask_supported(n * h * w, c) == 1
    g0 = torch.Generator(device=cuda_device).manual_seed(c)
    mk = lambda: torch.randn(shape, device=cuda_device, generator=g0).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last)
    x0, r0, dy = mk(), mk(), mk()
    gamma0 = (torch.rand(c, device=cuda_device, generator=g0) + 0.5).to(torch.bfloat16)
    beta0 = (torch.randn(c, device=cuda_device, generator=g0) * 0.3).to(torch.bfloat16)
    outs = []
    for use_mask in (True, False):
        monkeypatch.setattr(batchnorm, "_USE_MASK", use_mask)
        x, r = x0.clone().requires_grad_(True), r0.clone().requires_grad_(True)
        gamma, beta = gamma0.clone().requires_grad_(True), beta0.clone().requires_grad_(True)
        y = fused_bn_act(x, gamma, beta, None, None, r, True)
        saved = [t for t in y.grad_fn.saved_tensors if t is not None and t.dtype == torch.uint8]
        assert bool(saved) == use_mask
        if use_mask:        # the bits are the sign of the output
            bits = saved[0][: n * h * w * c // 8]
            want = (y.detach().permute(0, 2, 3, 1).reshape(-1, 8) > 0).to(torch.uint8)
            want = (want << torch.arange(8, device=cuda_device, dtype=torch.uint8)).sum(dim=1).to(torch.uint8)
            assert torch.equal(bits, want)
        y.backward(dy)
        outs.append((y.detach(), x.grad, r.grad, gamma.grad, beta.grad))
    for a, b in zip(*outs):
        assert torch.equal(a, b)


def test_module_train_and_eval_match_torch(cuda_device):
    torch.manual_seed(0)
    mod = FusedBatchNormAct(128, relu=True).to(cuda_device).to(torch.bfloat16)
    assert mod.running_mean.dtype == torch.float32          # statistics stay fp32
    ref = torch.nn.BatchNorm2d(128).to(cuda_device)
    x = torch.randn(4, 128, 12, 12, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    res = torch.randn_like(x)
    for _ in range(3):
        y = mod(x, res)
        yr = torch.relu(ref(x.float()) + res.float())
    assert torch.allclose(y.float(), yr, rtol=2e-2, atol=3e-2)
    assert torch.allclose(mod.running_mean, ref.running_mean, rtol=1e-3, atol=1e-3)
    assert torch.allclose(mod.running_var, ref.running_var, rtol=1e-3, atol=1e-3)
    mod.eval(); ref.eval()
    with torch.no_grad():
        ye = mod(x, res)
        yre = torch.relu(ref(x.float()) + res.float())
    assert torch.allclose(ye.float(), yre, rtol=2e-2, atol=3e-2)


def test_bn_apply_affine(cuda_device):
    x = torch.randn(2, 64, 8, 8, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    scale = torch.rand(64, device=cuda_device) + 0.5
    shift = torch.randn(64, device=cuda_device)
    y = bn_apply(x, scale, shift, None, False)
    want = x.float() * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)
    assert torch.allclose(y.float(), want, rtol=1e-2, atol=2e-2)


def test_resnet50_fused_vs_stock_forward(cuda_device):
    from tiresias_b200.models import vision
    torch.manual_seed(1)
    fused = vision.resnet50(fused=True).to(cuda_device).to(torch.bfloat16).to(memory_format=torch.channels_last)
    stock = vision.resnet50(fused=False).to(cuda_device).to(torch.bfloat16).to(memory_format=torch.channels_last)
    sd = fused.state_dict()
    remap = {}
    for k, v in sd.items():
        parts = k.split(".")
        if parts[-2].startswith("bn") or parts[-2].endswith("_bn"):
            parts.insert(-1, "bn")
        remap[".".join(parts)] = v
    remap = {k: (v.to(torch.bfloat16) if v.is_floating_point() else v) for k, v in remap.items()}
    missing = stock.load_state_dict(remap, strict=False)
    assert not [k for k in missing.missing_keys if "num_batches_tracked" not in k], missing
    x = torch.randn(8, 3, 96, 96, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    yf, ys = fused(x).float(), stock(x).float()
    assert torch.allclose(yf, ys, rtol=5e-2, atol=8e-2), (yf - ys).abs().max().item()


@pytest.mark.parametrize("shape", [(2, 64, 112, 112), (3, 16, 9, 7), (1, 8, 5, 6), (2, 24, 2, 2)])
def test_maxpool3x3s2_matches_torch_including_ties(cuda_device, shape):
    from tiresias_b200.ops.pooling import MaxPool3x3S2
    g = torch.Generator(device=cuda_device).manual_seed(sum(shape))
    # post-ReLU-like input: many exact zeros and (bf16) repeated values => plenty of ties
    x = torch.relu(torch.randn(shape, device=cuda_device, generator=g)).mul(4).round().div(4).to(torch.bfloat16)
    x = x.contiguous(memory_format=torch.channels_last).requires_grad_(True)
    y = MaxPool3x3S2()(x)
    dy = torch.randn(y.shape, device=cuda_device, generator=g).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y.backward(dy)
    xr = x.detach().float().requires_grad_(True)
    yr = torch.nn.functional.max_pool2d(xr, 3, 2, 1)
    yr.backward(dy.float())
    assert torch.equal(y.float(), yr)
    # torch accumulates several windows' gradients in fp32; ours in fp32 then rounds once to bf16
    assert torch.allclose(x.grad.float(), xr.grad, rtol=1e-2, atol=2e-2), (x.grad.float() - xr.grad).abs().max().item()
    assert abs(float(x.grad.float().sum()) - float(dy.float().sum())) < 0.05 * dy.numel() ** 0.5 + 1.0
