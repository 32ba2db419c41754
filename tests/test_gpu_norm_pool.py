"""LayerNorm (csrc/norm_kernels.cu) and the generic max-pool kernels (csrc/pool_kernels.cu) against plain PyTorch
fp32 references of the same ops."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,d", [(4096, 512), (8192, 1024), (1000, 256), (777, 768), (3, 512)])
@pytest.mark.parametrize("pdtype", [torch.bfloat16, torch.float32])
def test_layernorm_forward_backward(cuda_device, rows, d, pdtype):
    from tiresias_b200.ops.layernorm import FusedLayerNorm
    torch.manual_seed(rows + d)
    ln = FusedLayerNorm(d).to(cuda_device).to(pdtype)
    with torch.no_grad():
        ln.weight.copy_(torch.rand(d) + 0.5)
        ln.bias.copy_(torch.randn(d) * 0.3)
    x = (torch.randn(rows, d, device=cuda_device) * 1.5 + 0.3).to(torch.bfloat16).requires_grad_(True)
    dy = torch.randn(rows, d, device=cuda_device).to(torch.bfloat16)
    y = ln(x.view(1, rows, d)).view(rows, d)
    assert y.grad_fn is not None and type(y.grad_fn).__name__ != "NativeLayerNormBackward0"
    y.backward(dy)
    xf = x.detach().float().requires_grad_(True)
    wf = ln.weight.detach().float().requires_grad_(True)
    bf = ln.bias.detach().float().requires_grad_(True)
    yf = torch.nn.functional.layer_norm(xf, (d,), wf, bf, ln.eps)
    yf.backward(dy.float())
    assert torch.allclose(y.float(), yf, rtol=2e-2, atol=2e-2), (y.float() - yf).abs().max().item()
    assert torch.allclose(x.grad.float(), xf.grad, rtol=2e-2, atol=2e-2), (x.grad.float() - xf.grad).abs().max().item()
    tol = 2e-2 * max(1.0, rows ** 0.5)
    assert torch.allclose(ln.weight.grad.float(), wf.grad, rtol=2e-2, atol=tol)
    assert torch.allclose(ln.bias.grad.float(), bf.grad, rtol=2e-2, atol=tol)
    # deterministic
    g1 = ln.weight.grad.clone()
    ln.weight.grad = None
    x.grad = None
    ln(x.view(1, rows, d)).view(rows, d).backward(dy)
    assert torch.equal(ln.weight.grad, g1)


@pytest.mark.parametrize("k,s,p", [(3, 2, 1), (2, 2, 0), (3, 2, 0)])
@pytest.mark.parametrize("shape", [(2, 64, 56, 56), (3, 16, 13, 27), (1, 8, 7, 7), (2, 192, 27, 27)])
def test_maxpool_matches_torch(cuda_device, k, s, p, shape):
    from tiresias_b200.ops.pooling import MaxPool
    torch.manual_seed(k * 10 + s)
    # distinct values per window (ties are broken identically anyway: first maximum)
    x = torch.randn(shape, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xr = x.clone().requires_grad_(True)
    x.requires_grad_(True)
    y = MaxPool(k, s, p)(x)
    assert type(y.grad_fn).__name__ == "_MaxPoolFnBackward"
    yr = torch.nn.functional.max_pool2d(xr.float(), k, s, p)
    assert torch.equal(y.float(), yr)
    dy = torch.randn_like(yr)
    y.backward(dy.to(torch.bfloat16))
    yr.backward(dy.to(torch.bfloat16).float())
    # ties (equal bf16 values in one window) may route the gradient to a different element than torch's fp32 path
    bad = (x.grad.float() - xr.grad).abs() > 1e-2
    assert bad.float().mean().item() < 2e-3, bad.float().mean().item()
