"""Fused BatchNorm (+ residual add) (+ ReLU) for NHWC bf16 activations on the hand-written kernels of
csrc/bn_kernels.cu: forward = one statistics pass + one apply pass, backward = one reduction pass + one
apply pass that also emits the gradient of the residual branch. Replaces ``nn.BatchNorm2d`` + add + ReLU
in the ResNet job models (the reference names them only by checkpoint size, model/model_factory.py:47-55).

On CPU tensors (model wiring tests) the module falls back to ``F.batch_norm``; on CUDA tensors the native
kernels are mandatory.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
from torch import nn

from . import native

_scratch: Dict[tuple, torch.Tensor] = {}
_USE_MASK = os.environ.get("TB200_BN_MASK", "1") != "0"     # A/B switch for the ReLU bit mask (bench only)


def _get_scratch(device: torch.device) -> torch.Tensor:
    """One scratch (counters, partials, coefficients) per (device, stream): calls on one stream are ordered,
    and jobs that share a GPU (exec/ packs single-GPU jobs, each on its own streams) never share one."""
    dev = device.index if device.index is not None else torch.cuda.current_device()
    key = (dev, torch.cuda.current_stream(device).cuda_stream)
    buf = _scratch.get(key)
    if buf is None:
        buf = torch.zeros(native.BN_SCRATCH_FLOATS, dtype=torch.float32, device=device)
        _scratch[key] = buf
    return buf


def _rows_channels(x: torch.Tensor):
    """x is logically NCHW with channels_last strides (physically [N,H,W,C] contiguous)."""
    n, c, h, w = x.shape
    if not x.is_contiguous(memory_format=torch.channels_last):
        x = x.contiguous(memory_format=torch.channels_last)
    return x, n * h * w, c


class _FusedBNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, gamma, beta, running_mean, running_var, momentum, eps, relu, stats_partials=None):
        x, rows, c = _rows_channels(x)
        if residual is not None:
            residual, _, _ = _rows_channels(residual)
        y = torch.empty_like(x, memory_format=torch.channels_last)
        save_mean = torch.empty(c, dtype=torch.float32, device=x.device)
        save_invstd = torch.empty(c, dtype=torch.float32, device=x.device)
        scratch = _get_scratch(x.device)
        common = (x.data_ptr(), residual.data_ptr() if residual is not None else 0, y.data_ptr(),
                  gamma.data_ptr(), beta.data_ptr(), native.dtype_code(gamma.dtype),
                  running_mean.data_ptr() if running_mean is not None else 0,
                  running_var.data_ptr() if running_var is not None else 0,
                  float(momentum), float(eps), save_mean.data_ptr(), save_invstd.data_ptr(), scratch.data_ptr())
        # A residual + ReLU layer's backward needs the sign of every output: as one bit per element where the
        # streamed kernels serve the shape (the backward then reads 1/16 of a tensor instead of y, twice), else y
        mask = None
        if relu and residual is not None and _USE_MASK and native.lib().tb200_bn_mask_supported(rows, c):
            mask = torch.empty(rows * c // 8 + 16, dtype=torch.uint8, device=x.device)
        # stats_partials: the producing convolution's epilogue summed x already (fold + apply, no statistics pass)
        native.check(native.lib().tb200_bn_fwd_train_ex(
            *common, stats_partials.data_ptr() if stats_partials is not None else 0,
            stats_partials.shape[0] if stats_partials is not None else 0,
            mask.data_ptr() if mask is not None else 0, rows, c, int(relu), native.stream_ptr()))
        ctx.relu, ctx.has_res, ctx.rows, ctx.c = bool(relu), residual is not None, rows, c
        # the ReLU mask of a layer without residual input is recomputed from x in the backward
        ctx.save_for_backward(x, y if (relu and residual is not None and mask is None) else None, mask, gamma, beta,
                              save_mean, save_invstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, mask, gamma, beta, save_mean, save_invstd = ctx.saved_tensors
        dy, _, _ = _rows_channels(dy)
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        dres = torch.empty_like(x, memory_format=torch.channels_last) if ctx.has_res else None
        dgamma = torch.empty_like(gamma)
        dbeta = torch.empty_like(gamma)
        scratch = _get_scratch(x.device)
        native.check(native.lib().tb200_bn_bwd_ex(
            dy.data_ptr(), x.data_ptr(), y.data_ptr() if y is not None else 0,
            mask.data_ptr() if mask is not None else 0, gamma.data_ptr(),
            beta.data_ptr(), native.dtype_code(gamma.dtype), save_mean.data_ptr(), save_invstd.data_ptr(), dx.data_ptr(),
            dres.data_ptr() if dres is not None else 0, dgamma.data_ptr(), dbeta.data_ptr(),
            scratch.data_ptr(), ctx.rows, ctx.c, int(ctx.relu), native.stream_ptr()))
        return dx, dres, dgamma, dbeta, None, None, None, None, None, None


def fused_bn_act(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor,
                 running_mean: Optional[torch.Tensor], running_var: Optional[torch.Tensor],
                 residual: Optional[torch.Tensor] = None, relu: bool = False, momentum: float = 0.1,
                 eps: float = 1e-5, stats=None) -> torch.Tensor:
    """Training-mode ``relu?(batch_norm(x) + residual?)`` (bf16 NHWC, CUDA). ``stats``: the
    :class:`native.ConvStats` the convolution that produced ``x`` emitted from its epilogue (skips the
    statistics pass over ``x``)."""
    assert x.dtype == torch.bfloat16 and x.dim() == 4
    partials = None
    if stats is not None and stats.count == x.shape[0] * x.shape[2] * x.shape[3] and stats.partials.shape[1] == x.shape[1]:
        partials = stats.partials
    return _FusedBNFn.apply(x, residual, gamma, beta, running_mean, running_var, momentum, eps, relu, partials)


def bn_apply(x: torch.Tensor, scale: torch.Tensor, shift: torch.Tensor,
             residual: Optional[torch.Tensor] = None, relu: bool = False) -> torch.Tensor:
    """Inference: ``relu?(x*scale[c] + shift[c] + residual?)`` in one pass (no autograd)."""
    x, rows, c = _rows_channels(x)
    if residual is not None:
        residual, _, _ = _rows_channels(residual)
    y = torch.empty_like(x, memory_format=torch.channels_last)
    coef = torch.cat([scale.float().reshape(-1), shift.float().reshape(-1)]).contiguous()
    native.check(native.lib().tb200_bn_apply(x.data_ptr(), residual.data_ptr() if residual is not None else 0,
                                            y.data_ptr(), coef.data_ptr(), rows, c, int(relu),
                                            native.stream_ptr()))
    return y


class FusedBatchNormAct(nn.Module):
    """``BatchNorm2d`` with the following residual add and ReLU folded in.

    ``forward(x, residual=None)``. Affine parameters follow the model's dtype (bf16 inside the
    trainer's flat buffers); running statistics stay fp32 (``keep_fp32_buffers`` tells the trainer
    not to down-cast them)."""

    keep_fp32_buffers = True

    def __init__(self, num_features: int, relu: bool = False, eps: float = 1e-5, momentum: float = 0.1):
        super().__init__()
        if num_features % 8:
            raise ValueError("num_features must be a multiple of 8")
        self.num_features, self.relu, self.eps, self.momentum = num_features, relu, eps, momentum
        self.weight = nn.Parameter(torch.ones(num_features))
        self.bias = nn.Parameter(torch.zeros(num_features))
        self.register_buffer("running_mean", torch.zeros(num_features, dtype=torch.float32))
        self.register_buffer("running_var", torch.ones(num_features, dtype=torch.float32))

    def _apply(self, fn, recurse=True):
        # keep the running statistics in fp32 through .to(bfloat16) / .bfloat16()
        rm, rv = self.running_mean, self.running_var
        super()._apply(fn, recurse)
        for name, old in (("running_mean", rm), ("running_var", rv)):
            new = self._buffers[name]
            if new.dtype != torch.float32:
                self._buffers[name] = new.float()
        return self

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        if x.is_cuda and x.dtype == torch.bfloat16:
            if self.training:
                return fused_bn_act(x, self.weight, self.bias, self.running_mean, self.running_var,
                                    residual, self.relu, self.momentum, self.eps,
                                    stats=getattr(x, "_tb200_stats", None))
            invstd = torch.rsqrt(self.running_var + self.eps)
            scale = self.weight.float() * invstd
            return bn_apply(x, scale, self.bias.float() - self.running_mean * scale, residual, self.relu)
        y = torch.nn.functional.batch_norm(x.float(), self.running_mean, self.running_var, self.weight.float(),
                                           self.bias.float(), self.training, self.momentum, self.eps).to(x.dtype)
        if residual is not None:
            y = y + residual
        return torch.relu(y) if self.relu else y

    def extra_repr(self) -> str:
        return f"{self.num_features}, relu={self.relu}"
