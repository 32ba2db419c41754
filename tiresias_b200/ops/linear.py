"""Linear layers on the hand-written tcgen05 GEMM (K5).

Forward  : y  = act(x @ W^T + b)              one NT GEMM, bias/activation in the epilogue
Backward : dx = dy @ W                        W consumed as stored (MN-major B descriptor)
           dW = dy^T @ x                      dy and x consumed as stored (MN-major A and B)
           (widths that are not multiples of 8 fall back to explicit transposes + the NT GEMM)
The FC shapes that dominate the reference's model table (core/models.py:9-11: VGG fc6
25088x4096 = 392 MB fp32, fc7 4096x4096, fc8 4096x1000) all run through here.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import nn

from . import native


def _pad_rows(x: torch.Tensor, multiple: int) -> torch.Tensor:
    r = x.shape[0]
    pad = (-r) % multiple
    if pad == 0:
        return x
    return torch.cat([x, x.new_zeros((pad, x.shape[1]))], dim=0)


def _gemm_nt_pull(x2d: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], act: int,
                  pull_ctx) -> torch.Tensor:
    """K2: forward GEMM that gathers its row-sharded weight from the PS shard owners while it runs."""
    import ctypes as C
    trainer, bucket = pull_ctx
    m, k = x2d.shape
    n = weight.shape[0]
    y = torch.empty((m, n), dtype=torch.bfloat16, device=x2d.device)
    bias_ptr, bias_dt = (bias.data_ptr(), native.dtype_code(bias.dtype)) if bias is not None else (0, 0)
    native.check(native.lib().tb200_gemm_bf16_nt_pull(
        C.byref(trainer._world_struct), bucket.index, bucket.offset, bucket.pull_rows,
        x2d.data_ptr(), y.data_ptr(), bias_ptr, bias_dt, m, n, k, native.BF16, act,
        trainer.epoch.data_ptr(), trainer.status.data_ptr(), trainer.spin_budget, bucket.pull_sync.data_ptr(),
        native.stream_ptr()))
    return y


class _TcLinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x2d: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], act: int,
                pull_ctx=None):
        x2d = x2d.contiguous()
        if pull_ctx is not None:
            y = _gemm_nt_pull(x2d, weight, bias, act, pull_ctx)
        else:
            y = native.gemm_nt(x2d, weight, bias, act)
        ctx.act = act
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x2d, weight, y if act != native.ACT_NONE else None)
        return y

    @staticmethod
    def backward(ctx, dy: torch.Tensor):
        x2d, weight, y = ctx.saved_tensors
        dy = dy.contiguous()
        if ctx.act == native.ACT_RELU:
            dy = dy * (y > 0)
        elif ctx.act == native.ACT_GELU:
            raise NotImplementedError("fused GELU backward: use act=ACT_NONE + torch gelu")
        m = dy.shape[0]
        dx = dw = db = None
        n_out, k_in = weight.shape
        mn_ok = n_out % 8 == 0 and k_in % 8 == 0       # MN-major operands need 16-byte row strides
        if ctx.needs_input_grad[0]:
            if mn_ok:
                dx = native.gemm(dy, weight, b_mn_major=True)            # dy[M,N] . W[N,K], W read as stored
            else:
                dx = native.gemm_nt(dy, native.transpose(weight))
        if ctx.needs_input_grad[1]:
            if mn_ok:
                dw = native.gemm(dy, x2d, a_mn_major=True, b_mn_major=True)   # dy^T . x, no transposes
            else:
                dyt = native.transpose(_pad_rows(dy, 8))                  # [N,M']
                xt = native.transpose(_pad_rows(x2d, 8))                  # [K,M']
                dw = native.gemm_nt(dyt, xt)                              # [N,K]
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy.sum(dim=0, dtype=torch.float32).to(dy.dtype)
        return dx, dw, db, None, None


def tc_linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None,
              act: int = native.ACT_NONE, pull_ctx=None) -> torch.Tensor:
    lead = x.shape[:-1]
    y = _TcLinearFn.apply(x.reshape(-1, x.shape[-1]), weight, bias, act, pull_ctx)
    return y.reshape(*lead, weight.shape[0])


class TcLinear(nn.Module):
    """Drop-in ``nn.Linear`` (bf16) whose three GEMMs run on the tcgen05 kernel when the input
    lives on a CUDA device; on CPU (unit tests of model wiring) it uses ``F.linear``."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, act: int = native.ACT_NONE):
        super().__init__()
        if in_features % 8:
            raise ValueError("in_features must be a multiple of 8 (TMA 16-byte stride rule)")
        self.in_features, self.out_features, self.act = in_features, out_features, act
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.zeros(out_features)) if bias else None
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            # set by the trainer when this weight is row-sharded over the PS ranks (pull mode)
            pull_ctx = getattr(self, "_pull_ctx", None) if self.training else None
            return tc_linear(x, self.weight, self.bias, self.act, pull_ctx)
        y = torch.nn.functional.linear(x, self.weight, self.bias)
        if self.act == native.ACT_RELU:
            y = torch.relu(y)
        elif self.act == native.ACT_GELU:
            y = torch.nn.functional.gelu(y, approximate="tanh")
        return y

    def extra_repr(self) -> str:
        return f"{self.in_features}->{self.out_features}, act={self.act}"
