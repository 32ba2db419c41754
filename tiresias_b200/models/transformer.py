"""Transformer-base job model ("transformer" in model/model_factory.py:21-23; QKV/FFN shapes
512x{512,2048} per SURVEY.md §2.6 K5). Pre-LN stack, d_model 512, 8 heads, FFN 2048 with ReLU
(as in the original base model), tied token embedding, next-token loss on synthetic tokens.
All projections run on the tcgen05 GEMM, the LayerNorms on csrc/norm_kernels.cu and (head dimension 64) the
attention on the tcgen05 kernels of csrc/attention.cu; the stock arm (``fused=False``) uses torch throughout.
"""
from __future__ import annotations

import torch
from torch import nn
from torch.nn import functional as F

from ..ops import native
from ..ops.attention import attention_reference, fused_attention
from ..ops.layernorm import FusedLayerNorm
from ..ops.linear import TcLinear


class _TorchLinear(nn.Linear):
    """Stock nn.Linear (+ ReLU as a separate kernel): the PyTorch+NCCL baseline arm."""

    def __init__(self, cin: int, cout: int, bias: bool = True, relu: bool = False):
        super().__init__(cin, cout, bias=bias)
        self.relu = relu

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = super().forward(x)
        return torch.relu_(y) if self.relu else y


def _ln(d: int, fused: bool) -> nn.Module:
    return FusedLayerNorm(d) if fused else nn.LayerNorm(d)


def _lin(cin: int, cout: int, fused: bool, relu: bool = False, bias: bool = True) -> nn.Module:
    if fused:
        return TcLinear(cin, cout, bias=bias, act=native.ACT_RELU if relu else native.ACT_NONE)
    return _TorchLinear(cin, cout, bias=bias, relu=relu)


class Block(nn.Module):
    def __init__(self, d_model: int, n_heads: int, d_ff: int, fused: bool = True, causal: bool = True):
        super().__init__()
        self.n_heads, self.causal, self.fused = n_heads, causal, fused
        self.ln1 = _ln(d_model, fused)
        self.qkv = _lin(d_model, 3 * d_model, fused)
        self.proj = _lin(d_model, d_model, fused)
        self.ln2 = _ln(d_model, fused)
        self.ff1 = _lin(d_model, d_ff, fused, relu=True)
        self.ff2 = _lin(d_ff, d_model, fused)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        b, s, d = x.shape
        qkv = self.qkv(self.ln1(x)).view(b * s, 3 * d)
        attn = fused_attention if self.fused else attention_reference
        a = attn(qkv, b, s, self.n_heads, self.causal)
        x = x + self.proj(a.view(b, s, d))
        return x + self.ff2(self.ff1(self.ln2(x)))


class TransformerLM(nn.Module):
    def __init__(self, vocab: int = 32000, d_model: int = 512, n_layers: int = 12, n_heads: int = 8,
                 d_ff: int = 2048, max_len: int = 512, fused: bool = True):
        super().__init__()
        self.vocab = vocab
        self.tok = nn.Embedding(vocab, d_model)
        self.pos = nn.Embedding(max_len, d_model)
        self.blocks = nn.ModuleList(Block(d_model, n_heads, d_ff, fused) for _ in range(n_layers))
        self.ln_f = _ln(d_model, fused)
        self.head = _lin(d_model, vocab, fused, bias=False)
        nn.init.normal_(self.tok.weight, std=0.02)
        nn.init.normal_(self.pos.weight, std=0.02)

    def forward(self, tokens: torch.Tensor) -> torch.Tensor:
        b, s = tokens.shape
        x = self.tok(tokens) + self.pos(torch.arange(s, device=tokens.device))
        for blk in self.blocks:
            x = blk(x)
        return self.head(self.ln_f(x))


def transformer_base(vocab: int = 32000, fused: bool = True) -> TransformerLM:
    return TransformerLM(vocab=vocab, fused=fused)


def transformer_big(vocab: int = 32000, fused: bool = True) -> TransformerLM:
    """Sized towards the 1100 MB checkpoint of model_factory.py:23 (~275 M fp32 parameters)."""
    return TransformerLM(vocab=vocab, d_model=1024, n_layers=20, n_heads=16, d_ff=4096, fused=fused)
