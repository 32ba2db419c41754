"""Max pooling on the streaming kernels of csrc/pool_kernels.cu: 3x3 / stride 2 / pad 1 (the ResNet stem),
2x2 / stride 2 (VGG) and 3x3 / stride 2 / pad 0 (AlexNet, Inception). Drop-in for ``nn.MaxPool2d`` (floor mode)
on bf16 channels-last CUDA tensors; anything else falls back to torch."""
from __future__ import annotations

import torch
from torch import nn

from . import native


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, k: int, stride: int, pad: int) -> torch.Tensor:
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        n, c, h, w = x.shape
        oh, ow = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
        out = torch.empty((n, c, oh, ow), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        argmax = torch.empty((n, oh, ow, c), dtype=torch.uint8, device=x.device)     # kh*k+kw per element
        native.check(native.lib().tb200_maxpool_fwd(x.data_ptr(), out.data_ptr(), argmax.data_ptr(),
                                                   n, h, w, c, k, stride, pad, native.stream_ptr()))
        ctx.save_for_backward(argmax)
        ctx.geom = (n, c, h, w, k, stride, pad)
        return out

    @staticmethod
    def backward(ctx, dout: torch.Tensor):
        (argmax,) = ctx.saved_tensors
        if not dout.is_contiguous(memory_format=torch.channels_last):
            dout = dout.contiguous(memory_format=torch.channels_last)
        n, c, h, w, k, stride, pad = ctx.geom
        dx = torch.empty((n, c, h, w), dtype=dout.dtype, device=dout.device, memory_format=torch.channels_last)
        native.check(native.lib().tb200_maxpool_bwd(argmax.data_ptr(), dout.data_ptr(), dx.data_ptr(),
                                                   n, h, w, c, k, stride, pad, native.stream_ptr()))
        return dx, None, None, None


_SERVED = {(3, 2, 1), (2, 2, 0), (3, 2, 0)}


class MaxPool(nn.Module):
    """``nn.MaxPool2d(k, stride, pad)`` (floor mode). The native kernels serve ``(k, stride, pad)`` in
    {(3,2,1), (2,2,0), (3,2,0)} for bf16 channels-last CUDA inputs with a channel count that is a multiple of 8."""

    def __init__(self, kernel_size: int, stride: int, padding: int = 0):
        super().__init__()
        self.k, self.stride, self.pad = kernel_size, stride, padding

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and x.shape[1] % 8 == 0
                and (self.k, self.stride, self.pad) in _SERVED
                and x.shape[2] + 2 * self.pad >= self.k and x.shape[3] + 2 * self.pad >= self.k):
            return _MaxPoolFn.apply(x, self.k, self.stride, self.pad)
        return torch.nn.functional.max_pool2d(x, self.k, self.stride, self.pad)

    def extra_repr(self) -> str:
        return f"kernel_size={self.k}, stride={self.stride}, padding={self.pad}"


class MaxPool3x3S2(MaxPool):
    """The ResNet stem's pooling."""

    def __init__(self):
        super().__init__(3, 2, 1)
