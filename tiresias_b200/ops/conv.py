"""1x1 convolutions on the hand-written tcgen05 GEMM.

In NHWC a 1x1 / stride-1 convolution *is* a GEMM over the N*H*W pixel rows:
    y[R, Cout] = x[R, Cin] . W[Cout, Cin]^T          (forward, NT)
    dx[R, Cin] = dy[R, Cout] . W[Cout, Cin]          (W consumed as stored: MN-major B)
    dW[Cout, Cin] = dy[R, Cout]^T . x[R, Cin]        (dy and x consumed as stored: MN-major A and B)
so no im2col, no transposes and no layout change are needed — the activations are only re-viewed. Two
thirds of ResNet-50's convolutions (conv1 / conv3 of every bottleneck, the stride-1 projection) are of
this kind.

3x3 / stride 1 / pad 1 convolutions (``TcConv2d``) run forward and input gradient on the halo-tile kernel of
csrc/conv3x3.cu (the input rows a tile needs are fetched once, the nine taps are shifted shared-memory windows;
dgrad reads the forward weight as stored). Every other geometry is an implicit GEMM on the GEMM mainloop
(``tb200_conv2d_fprop`` / ``_dgrad`` / ``_wgrad``, csrc/gemm_tcgen05.cu): the producer warp gathers the activation
operand with im2col-mode TMA loads, so nothing is unfolded in memory; a strided input gradient runs as stride x
stride parity classes on that kernel. Forward passes can emit the per-channel sum / sum of squares of their output
from the epilogue; the fused BatchNorm that follows then skips its statistics pass (``emit_stats``).
"""
from __future__ import annotations

import math

import torch
from torch import nn

import os

from . import native

OWN_BIAS_CONV = os.environ.get("TB200_OWN_BIAS_CONV", "1") != "0"       # VGG / AlexNet 3x3 layers on our kernels
# widest image those layers take on our kernels (wider: segment-mode tiles and the im2col wgrad, measured slower than
# cuDNN on VGG-16's 224- and 112-pixel layers - profiles/bench_vgg16_r2f.md)
BIAS_CONV_MAX_W = int(os.environ.get("TB200_BIAS_CONV_MAX_W", "56"))
HALO_WGRAD = os.environ.get("TB200_HALO_WGRAD", "1") != "0"             # A/B switch (bench only)
FUSE_SKIP_GRAD = os.environ.get("TB200_FUSE_SKIP_GRAD", "1") != "0"     # A/B switch (bench only)


class GradBox:
    """Mailbox between the two consumers of a residual block's input: the skip path parks its gradient here
    (:class:`_ParkSkipGrad`) and the first convolution's dgrad GEMM adds it in its epilogue, so autograd has
    nothing left to sum (one element-wise add over the block input saved per block)."""

    __slots__ = ("grad", "main_done")

    def __init__(self):
        self.grad = None
        self.main_done = False


class _ParkSkipGrad(torch.autograd.Function):
    """Identity on the skip branch. Created AFTER the block's main branch in the forward, so its backward runs
    before the main branch's (autograd runs ready nodes latest-created first): the skip gradient is parked in the
    box and this node reports none. If the main branch's dgrad did run first, the gradient flows on as usual
    and autograd adds — the order only decides whether the add is fused."""

    @staticmethod
    def forward(ctx, x: torch.Tensor, box: GradBox):
        ctx.box = box
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        box = ctx.box
        if g is None or box.main_done:
            return g, None
        box.grad = g
        return None, None


def park_skip_grad(x: torch.Tensor, box: GradBox) -> torch.Tensor:
    return _ParkSkipGrad.apply(x, box)


class _Conv1x1Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, weight: torch.Tensor, ops: frozenset = frozenset("fdw"), want_stats: bool = False,
                box: GradBox = None):
        ctx.set_materialize_grads(False)      # no zeros kernel for the statistics output's (absent) gradient
        ctx.box = box
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        n, cin, h, w = x.shape
        cout = weight.shape[0]
        x2d = x.permute(0, 2, 3, 1).reshape(n * h * w, cin)               # a view of the NHWC storage
        w2d = weight.permute(0, 2, 3, 1).reshape(cout, cin)
        if not w2d.is_contiguous():
            w2d = w2d.contiguous()
        ctx.save_for_backward(x2d, w2d)
        ctx.shape = (n, cin, h, w, cout)
        ctx.ops = ops
        empty = x.new_empty(0, dtype=torch.float32)
        if "f" not in ops:
            ctx.mark_non_differentiable(empty)
            return torch.nn.functional.conv2d(x, w2d.view(cout, 1, 1, cin).permute(0, 3, 1, 2)), empty
        if want_stats:
            y2d, st = native.gemm_nt_stats(x2d, w2d)
            partials = st.partials if st is not None else empty
            ctx.mark_non_differentiable(partials)
            return y2d.view(n, h, w, cout).permute(0, 3, 1, 2), partials
        ctx.mark_non_differentiable(empty)
        y2d = native.gemm_nt(x2d, w2d)
        return y2d.view(n, h, w, cout).permute(0, 3, 1, 2), empty

    @staticmethod
    def backward(ctx, dy: torch.Tensor, _dstats=None):
        box = ctx.box
        skip = None
        if box is not None:                   # the skip branch's gradient, if its backward already ran
            skip, box.grad, box.main_done = box.grad, None, True
        if dy is None:
            return skip, None, None, None, None
        x2d, w2d = ctx.saved_tensors
        n, cin, h, w, cout = ctx.shape
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dy2d = dy.permute(0, 2, 3, 1).reshape(n * h * w, cout)
        ops = ctx.ops
        want_dx, want_dw = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        own_dx, own_dw = want_dx and "d" in ops, want_dw and "w" in ops
        dx = dw = None
        if own_dx:
            if skip is not None and skip.dtype == torch.bfloat16 and skip.shape == (n, cin, h, w):
                if not skip.is_contiguous(memory_format=torch.channels_last):
                    skip = skip.contiguous(memory_format=torch.channels_last)
                s2d = skip.permute(0, 2, 3, 1).reshape(n * h * w, cin)
                dx = native.gemm_add(dy2d, w2d, s2d, b_mn_major=True).view(n, h, w, cin).permute(0, 3, 1, 2)
                skip = None
            else:
                dx = native.gemm(dy2d, w2d, b_mn_major=True).view(n, h, w, cin).permute(0, 3, 1, 2)
        if own_dw:
            dw = native.gemm(dy2d, x2d, a_mn_major=True, b_mn_major=True).view(cout, 1, 1, cin).permute(0, 3, 1, 2)
        if (want_dx and not own_dx) or (want_dw and not own_dw):
            x4 = x2d.view(n, h, w, cin).permute(0, 3, 1, 2)
            w4 = w2d.view(cout, 1, 1, cin).permute(0, 3, 1, 2)
            lib_dx, lib_dw, _ = torch.ops.aten.convolution_backward(
                dy, x4, w4, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1,
                [want_dx and not own_dx, want_dw and not own_dw, False])
            dx = dx if own_dx else lib_dx
            dw = dw if own_dw else lib_dw
        if skip is not None:                  # parked but not fused (library dgrad, or no input gradient wanted)
            dx = skip if dx is None else dx + skip
        return dx, dw, None, None, None


def _attach_stats(y: torch.Tensor, partials: torch.Tensor) -> torch.Tensor:
    """Hang the epilogue's statistics on the activation so that the fused BatchNorm that consumes ``y`` next
    (``FusedBatchNormAct.forward``) skips its statistics pass."""
    if partials.numel():
        n, _, h, w = y.shape
        y._tb200_stats = native.ConvStats(partials, partials.shape[0], n * h * w)
    return y


class TcConv1x1(nn.Module):
    """``nn.Conv2d(cin, cout, 1, bias=False)`` whose three GEMMs run on the tcgen05 kernel for bf16
    channels-last CUDA inputs (anything else goes to ``F.conv2d``). The weight keeps Conv2d's shape
    ``[cout, cin, 1, 1]``, so checkpoints and the trainer's flat parameter layout are unchanged."""

    def __init__(self, in_channels: int, out_channels: int, ops=("f", "d", "w"), emit_stats: bool = False):
        super().__init__()
        self.emit_stats = emit_stats          # a fused BatchNorm follows: hand it the epilogue's statistics
        if in_channels % 8 or out_channels % 8:
            raise ValueError("channel counts must be multiples of 8 (TMA 16-byte stride rule)")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.ops = frozenset(ops)
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, 1, 1))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))

    def forward(self, x: torch.Tensor, grad_box: GradBox = None) -> torch.Tensor:
        """``grad_box``: see :class:`GradBox` (only honoured on the tcgen05 path; the caller checks
        :meth:`fuses_skip_grad` before parking anything)."""
        if x.is_cuda and x.dtype == torch.bfloat16 and self.weight.dtype == torch.bfloat16:
            want = self.emit_stats and self.training and torch.is_grad_enabled()
            return _attach_stats(*_Conv1x1Fn.apply(x, self.weight, self.ops, want, grad_box))
        return torch.nn.functional.conv2d(x, self.weight)

    def fuses_skip_grad(self, x: torch.Tensor) -> bool:
        return (x.is_cuda and x.dtype == torch.bfloat16 and self.weight.dtype == torch.bfloat16 and x.requires_grad
                and torch.is_grad_enabled() and FUSE_SKIP_GRAD)

    def extra_repr(self) -> str:
        return f"{self.in_channels}, {self.out_channels}, kernel_size=(1, 1), tcgen05"


def _parse_ops(spec: str):
    """``TB200_TC_CONV`` -> {kind: set of passes}. kind ``p`` = pointwise (1x1, stride 1), ``k`` = every
    other geometry; passes ``f`` forward, ``d`` input gradient, ``w`` weight gradient. ``1`` = everything,
    e.g. ``p:fw,k:f`` = pointwise forward + weight gradient and k x k forward on the tcgen05 kernels, the
    rest on cuDNN."""
    out = {"p": set(), "k": set()}
    spec = (spec or "").strip()
    if spec in ("", "0"):
        return out
    if spec == "1":
        return {"p": set("fdw"), "k": set("fdw")}
    for part in spec.split(","):
        kind, _, passes = part.partition(":")
        if kind not in out or not set(passes) <= set("fdw"):
            raise ValueError(f"bad TB200_TC_CONV entry {part!r}")
        out[kind] |= set(passes)
    return out


def _aten_backward(dy, x, weight, stride, pad, want_dx, want_dw):
    return torch.ops.aten.convolution_backward(dy, x, weight, None, [stride, stride], [pad, pad], [1, 1], False,
                                               [0, 0], 1, [want_dx, want_dw, False])


class _Conv2dFn(torch.autograd.Function):
    """One convolution whose three passes are individually routed to our tcgen05 kernels (``ops`` names them)
    or to cuDNN. 3x3 / stride 1 / pad 1 forward and input gradient run on the halo-tile kernel
    (csrc/conv3x3.cu); every other geometry on the im2col-TMA implicit GEMM (csrc/gemm_tcgen05.cu).
    Returns ``(y, stats_partials)``: with ``want_stats`` the forward's epilogue also emits the per-channel
    (sum, sum of squares) partials a following fused BatchNorm turns into its batch statistics."""

    @staticmethod
    def forward(ctx, x: torch.Tensor, weight: torch.Tensor, stride: int, pad: int, ops: frozenset, want_stats: bool):
        ctx.set_materialize_grads(False)
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        if not weight.is_contiguous(memory_format=torch.channels_last):
            weight = weight.contiguous(memory_format=torch.channels_last)
        ctx.save_for_backward(x, weight)
        ctx.geom = (stride, pad, ops)
        ctx.halo = weight.shape[2] == 3 and weight.shape[3] == 3 and stride == 1 and pad == 1 and x.shape[3] < 256
        empty = x.new_empty(0, dtype=torch.float32)
        ctx.nparts = 0
        if "f" in ops:
            if ctx.halo:
                if want_stats:
                    y, st = native.conv3x3_fprop(x, weight, stats=True)
                    ctx.nparts = st.nparts
                    ctx.mark_non_differentiable(st.partials)
                    return y, st.partials
            if want_stats:
                y, st = native.conv2d_fprop_stats(x, weight, stride, pad)
                partials = st.partials if st is not None else empty
                ctx.mark_non_differentiable(partials)
                return y, partials
            ctx.mark_non_differentiable(empty)
            if ctx.halo:
                return native.conv3x3_fprop(x, weight), empty
            return native.conv2d_fprop(x, weight, stride, pad), empty
        ctx.mark_non_differentiable(empty)
        return torch.nn.functional.conv2d(x, weight, stride=stride, padding=pad), empty

    @staticmethod
    def backward(ctx, dy: torch.Tensor, _dstats=None):
        if dy is None:
            return None, None, None, None, None, None
        x, weight = ctx.saved_tensors
        stride, pad, ops = ctx.geom
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        want_dx, want_dw = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        # (a strided input gradient runs as stride x stride parity classes on the same im2col kernel)
        own_dx = want_dx and "d" in ops and weight.shape[0] % 64 == 0
        own_dw = want_dw and "w" in ops
        dx = dw = None
        if own_dx:
            dx = native.conv3x3_dgrad(dy, weight) if ctx.halo else native.conv2d_dgrad(dy, weight, x.shape, stride, pad)
        if own_dw:
            if ctx.halo and HALO_WGRAD and native.conv3x3_wgrad_eligible(x, weight.shape[0]):
                dw = native.conv3x3_wgrad(x, dy)
            else:
                dw = native.conv2d_wgrad(x, dy, weight.shape[2:], stride, pad)
        if (want_dx and not own_dx) or (want_dw and not own_dw):
            lib_dx, lib_dw, _ = _aten_backward(dy, x, weight, stride, pad, want_dx and not own_dx,
                                               want_dw and not own_dw)
            dx = dx if own_dx else lib_dx
            dw = dw if own_dw else lib_dw
        return dx, dw, None, None, None, None


class TcConv2d(nn.Module):
    """``nn.Conv2d(cin, cout, k, stride, padding, bias=False)`` as an implicit GEMM on the tcgen05 kernel
    for bf16 channels-last CUDA inputs (anything else goes to ``F.conv2d``). ``cin`` must be a multiple of
    64 (one im2col box = 64 channels of one filter tap) and ``cout`` of 8. Same parameter shape as
    Conv2d. ``ops`` picks the passes that run on our kernels (default: all three)."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size: int, stride: int = 1,
                 padding: int = 0, ops=("f", "d", "w"), emit_stats: bool = False):
        super().__init__()
        self.emit_stats = emit_stats          # a fused BatchNorm follows: hand it the epilogue's statistics
        if in_channels % 64 or out_channels % 8:
            raise ValueError("TcConv2d needs in_channels % 64 == 0 and out_channels % 8 == 0")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.padding = kernel_size, stride, padding
        self.ops = frozenset(ops)
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, kernel_size, kernel_size))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda and x.dtype == torch.bfloat16 and self.weight.dtype == torch.bfloat16:
            want = self.emit_stats and self.training and torch.is_grad_enabled()
            return _attach_stats(*_Conv2dFn.apply(x, self.weight, self.stride, self.padding, self.ops, want))
        return torch.nn.functional.conv2d(x, self.weight, stride=self.stride, padding=self.padding)

    def extra_repr(self) -> str:
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, "
                f"stride={self.stride}, padding={self.padding}, tcgen05 implicit GEMM passes={sorted(self.ops)}")


class _Conv3x3BiasActFn(torch.autograd.Function):
    """``relu?(conv3x3(x, w) + b)`` for the VGG / AlexNet style layers: bias and ReLU in the halo-tile kernel's epilogue;
    backward = ReLU mask on dy (element-wise), halo-tile dgrad and wgrad (im2col wgrad where the halo one does not
    serve the geometry), bias gradient as a channel sum."""

    @staticmethod
    def forward(ctx, x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, relu: bool):
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        if not weight.is_contiguous(memory_format=torch.channels_last):
            weight = weight.contiguous(memory_format=torch.channels_last)
        y = native.conv3x3_fprop_bias(x, weight, bias, relu)
        ctx.save_for_backward(x, weight, y if relu else None)
        ctx.relu = relu
        return y

    @staticmethod
    def backward(ctx, dy: torch.Tensor):
        x, weight, y = ctx.saved_tensors
        if ctx.relu:
            dy = dy * (y > 0)
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = native.conv3x3_dgrad(dy, weight)
        if ctx.needs_input_grad[1]:
            if HALO_WGRAD and native.conv3x3_wgrad_eligible(x, weight.shape[0]):
                dw = native.conv3x3_wgrad(x, dy)
            else:
                dw = native.conv2d_wgrad(x, dy, (3, 3), 1, 1)
        if ctx.needs_input_grad[2]:
            db = dy.sum(dim=(0, 2, 3), dtype=torch.float32).to(dy.dtype)
        return dx, dw, db, None


class TcConv3x3Bias(nn.Module):
    """``nn.Conv2d(cin, cout, 3, 1, 1)`` (with bias) + optional ReLU on the halo-tile kernels. ``cin`` and ``cout``
    multiples of 64. Same parameter names and shapes as Conv2d, so checkpoints and the flat parameter layout match the
    stock network's; anything the kernels do not serve goes to ``F.conv2d`` (+ ``relu``)."""

    def __init__(self, in_channels: int, out_channels: int, relu: bool = True):
        super().__init__()
        if in_channels % 64 or out_channels % 64:
            raise ValueError("TcConv3x3Bias needs channel counts that are multiples of 64")
        self.in_channels, self.out_channels, self.relu = in_channels, out_channels, relu
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, 3, 3))
        self.bias = nn.Parameter(torch.zeros(out_channels))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        bound = 1.0 / math.sqrt(in_channels * 9)
        nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if (x.is_cuda and x.dtype == torch.bfloat16 and self.weight.dtype == torch.bfloat16 and OWN_BIAS_CONV
                and x.shape[3] <= BIAS_CONV_MAX_W
                and native.conv3x3_supported(x.shape[0], x.shape[2], x.shape[3], self.in_channels, self.out_channels)
                and native.conv3x3_supported(x.shape[0], x.shape[2], x.shape[3], self.out_channels, self.in_channels, True)):
            return _Conv3x3BiasActFn.apply(x, self.weight, self.bias, self.relu)
        y = torch.nn.functional.conv2d(x, self.weight, self.bias, stride=1, padding=1)
        return torch.relu(y) if self.relu else y

    def extra_repr(self) -> str:
        return f"{self.in_channels}, {self.out_channels}, kernel_size=(3, 3), bias, relu={self.relu}, tcgen05 halo tile"


class _PackedStemFn(torch.autograd.Function):
    """The ResNet stem (7x7 / stride 2 / pad 3 over 3 channels) on the packed input of csrc/stem.cu: a 4x1 /
    stride-1 implicit GEMM with 64 input channels. ``y_packed`` [N,64,OH+3,OW] comes from
    ``native.stem7x7_pack`` (the image's uint8 -> bf16 normalisation is fused into that pack) and needs no
    gradient; the weight keeps its [cout,3,7,7] shape."""

    @staticmethod
    def forward(ctx, y_packed: torch.Tensor, weight: torch.Tensor, want_stats: bool):
        ctx.set_materialize_grads(False)
        if not weight.is_contiguous(memory_format=torch.channels_last):
            weight = weight.contiguous(memory_format=torch.channels_last)
        wp = native.stem7x7_weight_pack(weight)
        ctx.save_for_backward(y_packed)
        ctx.wdtype = weight.dtype
        empty = y_packed.new_empty(0, dtype=torch.float32)
        if want_stats:
            y, st = native.conv2d_fprop_stats(y_packed, wp, 1, 0)
            partials = st.partials if st is not None else empty
            ctx.mark_non_differentiable(partials)
            return y, partials
        ctx.mark_non_differentiable(empty)
        return native.conv2d_fprop(y_packed, wp, 1, 0), empty

    @staticmethod
    def backward(ctx, dy: torch.Tensor, _dstats=None):
        if dy is None:
            return None, None, None
        (y_packed,) = ctx.saved_tensors
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dwp = native.conv2d_wgrad(y_packed, dy, (4, 1), 1, 0, out_dtype=torch.float32)
        return None, native.stem7x7_wgrad_unpack(dwp, ctx.wdtype), None


def packed_stem_conv(y_packed: torch.Tensor, weight: torch.Tensor, emit_stats: bool = True) -> torch.Tensor:
    want = emit_stats and torch.is_grad_enabled() and weight.requires_grad
    return _attach_stats(*_PackedStemFn.apply(y_packed, weight, want))
