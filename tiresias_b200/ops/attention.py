"""Multi-head attention (head dimension 64) on the tcgen05 kernels of csrc/attention.cu.

``fused_attention(qkv, batch, seq, heads, causal)`` takes the fused projection output ``[batch*seq, 3*heads*64]``
(columns Q | K | V) and returns ``[batch*seq, heads*64]``: the kernels read the heads in place and the backward
writes the gradient in the same fused layout, so there are no per-head transposes around the attention.
Anything the kernels do not serve (CPU tensors, other head sizes, fp32) goes through torch SDPA
(:func:`attention_reference`, also the numerics oracle of the GPU test). The Transformer / BERT job models are the
users (reference: model/model_factory.py:19-23)."""
from __future__ import annotations

import os

import torch
from torch.nn import functional as F

from . import native

OWN_ATTENTION = os.environ.get("TB200_OWN_ATTENTION", "1") != "0"      # A/B switch (bench only)


def attention_reference(qkv: torch.Tensor, batch: int, seq: int, heads: int, causal: bool) -> torch.Tensor:
    d = qkv.shape[-1] // 3
    q, k, v = (t.transpose(1, 2) for t in qkv.view(batch, seq, 3, heads, d // heads).unbind(dim=2))
    a = F.scaled_dot_product_attention(q, k, v, is_causal=causal)
    return a.transpose(1, 2).reshape(batch * seq, d)


class _AttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv: torch.Tensor, batch: int, seq: int, heads: int, causal: bool):
        qkv = qkv.contiguous()
        out = torch.empty((batch * seq, heads * 64), dtype=torch.bfloat16, device=qkv.device)
        lse2 = torch.empty((batch, heads, seq), dtype=torch.float32, device=qkv.device)
        native.check(native.lib().tb200_attention_fwd(qkv.data_ptr(), out.data_ptr(), lse2.data_ptr(), batch, seq, heads,
                                                     int(causal), native.stream_ptr()))
        ctx.save_for_backward(qkv, out, lse2)
        ctx.geom = (batch, seq, heads, causal)
        return out

    @staticmethod
    def backward(ctx, dout: torch.Tensor):
        qkv, out, lse2 = ctx.saved_tensors
        batch, seq, heads, causal = ctx.geom
        dout = dout.contiguous()
        dqkv = torch.empty_like(qkv)
        delta = torch.empty_like(lse2)
        native.check(native.lib().tb200_attention_bwd(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse2.data_ptr(),
                                                     delta.data_ptr(), dqkv.data_ptr(), batch, seq, heads, int(causal),
                                                     native.stream_ptr()))
        return dqkv, None, None, None, None


def served(qkv: torch.Tensor, heads: int) -> bool:
    return (OWN_ATTENTION and qkv.is_cuda and qkv.dtype == torch.bfloat16 and qkv.dim() == 2
            and qkv.shape[1] == 3 * heads * 64)


def fused_attention(qkv: torch.Tensor, batch: int, seq: int, heads: int, causal: bool) -> torch.Tensor:
    """``qkv`` [batch*seq, 3*heads*head_dim] -> attention output [batch*seq, heads*head_dim]."""
    if served(qkv, heads):
        return _AttentionFn.apply(qkv, batch, seq, heads, causal)
    return attention_reference(qkv, batch, seq, heads, causal)
