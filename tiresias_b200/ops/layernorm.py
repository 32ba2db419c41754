"""LayerNorm over the last dimension on the kernels of csrc/norm_kernels.cu (one warp per row, the row in
registers; deterministic dgamma / dbeta). Drop-in for ``nn.LayerNorm(d)`` on bf16 CUDA inputs with d in
{256, 512, 768, 1024}; anything else (CPU tests, other widths, fp32) goes to ``F.layer_norm``.
The Transformer / BERT job models (reference: model/model_factory.py:19-23) normalise every block input."""
from __future__ import annotations

from typing import Dict

import torch
from torch import nn

from . import native

_scratch: Dict[tuple, torch.Tensor] = {}


def _get_scratch(device: torch.device, d: int) -> torch.Tensor:
    dev = device.index if device.index is not None else torch.cuda.current_device()
    key = (dev, torch.cuda.current_stream(device).cuda_stream, d)
    buf = _scratch.get(key)
    if buf is None:
        buf = torch.empty(native.lib().tb200_layernorm_scratch_floats(d), dtype=torch.float32, device=device)
        _scratch[key] = buf
    return buf


class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float):
        shape = x.shape
        d = shape[-1]
        x2 = x.reshape(-1, d)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        native.check(native.lib().tb200_layernorm_fwd(x2.data_ptr(), y.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
                                                     native.dtype_code(gamma.dtype), mean.data_ptr(), rstd.data_ptr(),
                                                     rows, d, float(eps), native.stream_ptr()))
        ctx.save_for_backward(x2, gamma, mean, rstd)
        ctx.shape = shape
        return y.view(shape)

    @staticmethod
    def backward(ctx, dy: torch.Tensor):
        x2, gamma, mean, rstd = ctx.saved_tensors
        d = x2.shape[1]
        dy2 = dy.reshape(-1, d)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = torch.empty_like(x2)
        dgamma = torch.empty_like(gamma)
        dbeta = torch.empty_like(gamma)
        scratch = _get_scratch(x2.device, d)
        native.check(native.lib().tb200_layernorm_bwd(dy2.data_ptr(), x2.data_ptr(), gamma.data_ptr(),
                                                     native.dtype_code(gamma.dtype), mean.data_ptr(), rstd.data_ptr(),
                                                     dx.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), scratch.data_ptr(),
                                                     scratch.numel(), x2.shape[0], d, native.stream_ptr()))
        return dx.view(ctx.shape), dgamma, dbeta, None


class FusedLayerNorm(nn.LayerNorm):
    """``nn.LayerNorm(d)`` whose forward / backward run on the native kernels when they serve the input."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        d = x.shape[-1]
        if (x.is_cuda and x.dtype == torch.bfloat16 and self.elementwise_affine and self.bias is not None
                and len(self.normalized_shape) == 1 and d in (256, 512, 768, 1024)
                and self.weight.dtype in (torch.bfloat16, torch.float32) and self.weight.dtype == self.bias.dtype):
            return _LayerNormFn.apply(x, self.weight, self.bias, self.eps)
        return super().forward(x)
