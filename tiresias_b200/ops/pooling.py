"""3x3 / stride 2 / pad 1 max pooling (the ResNet stem) on the streaming kernels of csrc/pool_kernels.cu.
Drop-in for ``nn.MaxPool2d(3, 2, 1)`` on bf16 channels-last CUDA tensors; anything else falls back to torch."""
from __future__ import annotations

import torch
from torch import nn

from . import native


class _MaxPool3x3S2Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor) -> torch.Tensor:
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        n, c, h, w = x.shape
        oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
        out = torch.empty((n, c, oh, ow), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        argmax = torch.empty((n, oh, ow, c), dtype=torch.uint8, device=x.device)     # kh*3+kw per element
        native.check(native.lib().tb200_maxpool3x3s2_fwd(x.data_ptr(), out.data_ptr(), argmax.data_ptr(),
                                                        n, h, w, c, native.stream_ptr()))
        ctx.save_for_backward(argmax)
        ctx.in_shape = (n, c, h, w)
        return out

    @staticmethod
    def backward(ctx, dout: torch.Tensor):
        (argmax,) = ctx.saved_tensors
        if not dout.is_contiguous(memory_format=torch.channels_last):
            dout = dout.contiguous(memory_format=torch.channels_last)
        n, c, h, w = ctx.in_shape
        dx = torch.empty((n, c, h, w), dtype=dout.dtype, device=dout.device, memory_format=torch.channels_last)
        native.check(native.lib().tb200_maxpool3x3s2_bwd(argmax.data_ptr(), dout.data_ptr(), dx.data_ptr(),
                                                        n, h, w, c, native.stream_ptr()))
        return dx


class MaxPool3x3S2(nn.Module):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and x.shape[1] % 8 == 0:
            return _MaxPool3x3S2Fn.apply(x)
        return torch.nn.functional.max_pool2d(x, 3, 2, 1)
