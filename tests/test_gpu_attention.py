"""tcgen05 attention (csrc/attention.cu) against an fp32 PyTorch reference of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _reference(qkv, b, s, h, causal):
    q, k, v = (t.transpose(1, 2) for t in qkv.view(b, s, 3, h, 64).unbind(dim=2))
    att = (q @ k.transpose(-1, -2)) * 0.125
    if causal:
        att = att.masked_fill(torch.triu(torch.ones(s, s, dtype=torch.bool, device=qkv.device), 1), float("-inf"))
    return (torch.softmax(att, dim=-1) @ v).transpose(1, 2).reshape(b * s, h * 64)


@pytest.mark.parametrize("b,s,h", [(2, 128, 2), (3, 256, 4), (2, 100, 3), (1, 384, 1), (4, 64, 8), (2, 130, 2)])
@pytest.mark.parametrize("causal", [True, False])
def test_attention_forward_backward(cuda_device, b, s, h, causal):
    from tiresias_b200.ops.attention import _AttentionFn, served
    torch.manual_seed(b * 1000 + s + h)
    qkv = (torch.randn(b * s, 3 * h * 64, device=cuda_device) * 0.8).to(torch.bfloat16).requires_grad_(True)
    assert served(qkv, h)
    out = _AttentionFn.apply(qkv, b, s, h, causal)
    dout = torch.randn(b * s, h * 64, device=cuda_device).to(torch.bfloat16)
    out.backward(dout)
    ref_in = qkv.detach().float().requires_grad_(True)
    ref = _reference(ref_in, b, s, h, causal)
    ref.backward(dout.float())
    assert torch.allclose(out.float(), ref, rtol=2e-2, atol=2e-2), (out.float() - ref).abs().max().item()
    err = (qkv.grad.float() - ref_in.grad).abs()
    scale = ref_in.grad.abs().max().item()
    assert err.max().item() <= 3e-2 * max(scale, 1.0), (err.max().item(), scale)
    # deterministic (no atomics anywhere)
    g1 = qkv.grad.clone()
    qkv.grad = None
    _AttentionFn.apply(qkv, b, s, h, causal).backward(dout)
    assert torch.equal(qkv.grad, g1)


def test_transformer_block_uses_own_attention_and_matches_stock(cuda_device):
    from tiresias_b200.models.transformer import Block
    torch.manual_seed(0)
    fused = Block(512, 8, 2048, fused=True).to(cuda_device).to(torch.bfloat16)
    stock = Block(512, 8, 2048, fused=False).to(cuda_device).to(torch.bfloat16)
    stock.load_state_dict(fused.state_dict())
    x = torch.randn(4, 128, 512, device=cuda_device).to(torch.bfloat16)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ya, yb = fused(xa), stock(xb)
    assert torch.allclose(ya.float(), yb.float(), rtol=5e-2, atol=5e-2), (ya.float() - yb.float()).abs().max().item()
    g = torch.randn_like(ya)
    ya.backward(g); yb.backward(g)
    assert torch.allclose(xa.grad.float(), xb.grad.float(), rtol=5e-2, atol=5e-2 * xb.grad.float().abs().max().item())
