"""GoogLeNet (Inception v1 with BatchNorm), Inception v3 and MobileNet v1 (width 0.25): the real architectures
behind the trace's ``googlenet`` (26 MB), ``inception3`` (104 MB) and ``mobilenet_v1_025`` checkpoints
(model/model_factory.py:31-45), so that a job of that name moves that model's gradient volume — round 1 ran
ResNet-50 in their place, up to 6x too many bytes. Random init, NHWC, bf16.

Every convolution is followed by a BatchNorm + ReLU, which run on the fused kernels
(:class:`tiresias_b200.ops.batchnorm.FusedBatchNormAct`; all channel counts here are multiples of 8). 1x1 and
3x3 / stride 1 / pad 1 convolutions with tensor-core-friendly channel counts go to our tcgen05 kernels through
``vision._conv``; the odd geometries of these networks (1x7, 7x1, 5x5, depthwise, unpadded stride 2) stay on
cuDNN. ``fused=False`` builds the same graph from stock PyTorch layers only (baseline arm).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple, Union

import torch
from torch import nn
from torch.nn import functional as F

from . import vision

_Size = Union[int, Tuple[int, int]]


def _pair(v: _Size) -> Tuple[int, int]:
    return (v, v) if isinstance(v, int) else v


class ConvBN(nn.Module):
    """conv (no bias) -> BatchNorm -> ReLU."""

    def __init__(self, cin: int, cout: int, k: _Size, stride: int = 1, pad: _Size = 0, groups: int = 1,
                 fused: bool = True):
        super().__init__()
        (kh, kw), (ph, pw) = _pair(k), _pair(pad)
        if groups == 1 and kh == kw and ph == pw:
            self.conv = vision._conv(cin, cout, kh, stride, ph, fused=fused)
        else:
            self.conv = nn.Conv2d(cin, cout, (kh, kw), stride, (ph, pw), groups=groups, bias=False)
        self.bn = vision._bn(cout, True, fused)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.bn(self.conv(x))


def _cat(xs: Sequence[torch.Tensor]) -> torch.Tensor:
    return torch.cat(list(xs), 1).contiguous(memory_format=torch.channels_last)


# ---- GoogLeNet ------------------------------------------------------------------------------------------------
class InceptionV1(nn.Module):
    def __init__(self, cin: int, c1: int, c3r: int, c3: int, c5r: int, c5: int, cp: int, fused: bool):
        super().__init__()
        self.b1 = ConvBN(cin, c1, 1, fused=fused)
        self.b2 = nn.Sequential(ConvBN(cin, c3r, 1, fused=fused), ConvBN(c3r, c3, 3, pad=1, fused=fused))
        self.b3 = nn.Sequential(ConvBN(cin, c5r, 1, fused=fused), ConvBN(c5r, c5, 3, pad=1, fused=fused))
        self.b4 = nn.Sequential(nn.MaxPool2d(3, 1, 1), ConvBN(cin, cp, 1, fused=fused))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _cat((self.b1(x), self.b2(x), self.b3(x), self.b4(x)))


class GoogLeNet(nn.Module):
    """~6.6 M parameters (26 MB fp32, the table's ``googlenet``: 26)."""

    def __init__(self, num_classes: int = 1000, fused: bool = True):
        super().__init__()
        f = fused
        self.stem = nn.Sequential(ConvBN(3, 64, 7, 2, 3, fused=f), nn.MaxPool2d(3, 2, ceil_mode=True),
                                  ConvBN(64, 64, 1, fused=f), ConvBN(64, 192, 3, pad=1, fused=f),
                                  nn.MaxPool2d(3, 2, ceil_mode=True))
        cfg = [(192, 64, 96, 128, 16, 32, 32), (256, 128, 128, 192, 32, 96, 64), "M",
               (480, 192, 96, 208, 16, 48, 64), (512, 160, 112, 224, 24, 64, 64), (512, 128, 128, 256, 24, 64, 64),
               (512, 112, 144, 288, 32, 64, 64), (528, 256, 160, 320, 32, 128, 128), "M",
               (832, 256, 160, 320, 32, 128, 128), (832, 384, 192, 384, 48, 128, 128)]
        layers: List[nn.Module] = []
        for c in cfg:
            layers.append(nn.MaxPool2d(3, 2, ceil_mode=True) if c == "M" else InceptionV1(*c, fused=f))
        self.blocks = nn.Sequential(*layers)
        self.fc = vision.TcLinear(1024, num_classes) if f else nn.Linear(1024, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.blocks(self.stem(x))
        return self.fc(torch.flatten(F.adaptive_avg_pool2d(x, 1), 1))


# ---- Inception v3 -----------------------------------------------------------------------------------------------
class _IncA(nn.Module):
    def __init__(self, cin: int, pool: int, f: bool):
        super().__init__()
        self.b1 = ConvBN(cin, 64, 1, fused=f)
        self.b5 = nn.Sequential(ConvBN(cin, 48, 1, fused=f), ConvBN(48, 64, 5, pad=2, fused=f))
        self.b3 = nn.Sequential(ConvBN(cin, 64, 1, fused=f), ConvBN(64, 96, 3, pad=1, fused=f),
                                ConvBN(96, 96, 3, pad=1, fused=f))
        self.bp = ConvBN(cin, pool, 1, fused=f)

    def forward(self, x):
        return _cat((self.b1(x), self.b5(x), self.b3(x), self.bp(F.avg_pool2d(x, 3, 1, 1))))


class _IncB(nn.Module):
    def __init__(self, cin: int, f: bool):
        super().__init__()
        self.b3 = ConvBN(cin, 384, 3, 2, fused=f)
        self.bd = nn.Sequential(ConvBN(cin, 64, 1, fused=f), ConvBN(64, 96, 3, pad=1, fused=f),
                                ConvBN(96, 96, 3, 2, fused=f))

    def forward(self, x):
        return _cat((self.b3(x), self.bd(x), F.max_pool2d(x, 3, 2)))


class _IncC(nn.Module):
    def __init__(self, cin: int, c7: int, f: bool):
        super().__init__()
        self.b1 = ConvBN(cin, 192, 1, fused=f)
        self.b7 = nn.Sequential(ConvBN(cin, c7, 1, fused=f), ConvBN(c7, c7, (1, 7), pad=(0, 3), fused=f),
                                ConvBN(c7, 192, (7, 1), pad=(3, 0), fused=f))
        self.bd = nn.Sequential(ConvBN(cin, c7, 1, fused=f), ConvBN(c7, c7, (7, 1), pad=(3, 0), fused=f),
                                ConvBN(c7, c7, (1, 7), pad=(0, 3), fused=f), ConvBN(c7, c7, (7, 1), pad=(3, 0), fused=f),
                                ConvBN(c7, 192, (1, 7), pad=(0, 3), fused=f))
        self.bp = ConvBN(cin, 192, 1, fused=f)

    def forward(self, x):
        return _cat((self.b1(x), self.b7(x), self.bd(x), self.bp(F.avg_pool2d(x, 3, 1, 1))))


class _IncD(nn.Module):
    def __init__(self, cin: int, f: bool):
        super().__init__()
        self.b3 = nn.Sequential(ConvBN(cin, 192, 1, fused=f), ConvBN(192, 320, 3, 2, fused=f))
        self.b7 = nn.Sequential(ConvBN(cin, 192, 1, fused=f), ConvBN(192, 192, (1, 7), pad=(0, 3), fused=f),
                                ConvBN(192, 192, (7, 1), pad=(3, 0), fused=f), ConvBN(192, 192, 3, 2, fused=f))

    def forward(self, x):
        return _cat((self.b3(x), self.b7(x), F.max_pool2d(x, 3, 2)))


class _IncE(nn.Module):
    def __init__(self, cin: int, f: bool):
        super().__init__()
        self.b1 = ConvBN(cin, 320, 1, fused=f)
        self.b3 = ConvBN(cin, 384, 1, fused=f)
        self.b3a = ConvBN(384, 384, (1, 3), pad=(0, 1), fused=f)
        self.b3b = ConvBN(384, 384, (3, 1), pad=(1, 0), fused=f)
        self.bd = nn.Sequential(ConvBN(cin, 448, 1, fused=f), ConvBN(448, 384, 3, pad=1, fused=f))
        self.bda = ConvBN(384, 384, (1, 3), pad=(0, 1), fused=f)
        self.bdb = ConvBN(384, 384, (3, 1), pad=(1, 0), fused=f)
        self.bp = ConvBN(cin, 192, 1, fused=f)

    def forward(self, x):
        b3 = self.b3(x)
        bd = self.bd(x)
        return _cat((self.b1(x), self.b3a(b3), self.b3b(b3), self.bda(bd), self.bdb(bd),
                     self.bp(F.avg_pool2d(x, 3, 1, 1))))


class InceptionV3(nn.Module):
    """~23.8 M parameters without the auxiliary head (95 MB fp32; with it the table's ``inception3``: 104).
    Runs at 299x299 in the wild; any input from 75x75 up works (the trace jobs feed 224x224)."""

    def __init__(self, num_classes: int = 1000, fused: bool = True, aux: bool = True):
        super().__init__()
        f = fused
        self.stem = nn.Sequential(ConvBN(3, 32, 3, 2, fused=f), ConvBN(32, 32, 3, fused=f), ConvBN(32, 64, 3, pad=1, fused=f),
                                  nn.MaxPool2d(3, 2), ConvBN(64, 80, 1, fused=f), ConvBN(80, 192, 3, fused=f),
                                  nn.MaxPool2d(3, 2))
        self.mixed = nn.Sequential(_IncA(192, 32, f), _IncA(256, 64, f), _IncA(288, 64, f), _IncB(288, f),
                                   _IncC(768, 128, f), _IncC(768, 160, f), _IncC(768, 160, f), _IncC(768, 192, f))
        self.tail = nn.Sequential(_IncD(768, f), _IncE(1280, f), _IncE(2048, f))
        # auxiliary classifier: part of the checkpoint (and of the gradient volume) in training
        self.aux = nn.Sequential(ConvBN(768, 128, 1, fused=f), ConvBN(128, 768, 5, pad=2, fused=f)) if aux else None
        self.aux_fc = (vision.TcLinear(768, num_classes) if f else nn.Linear(768, num_classes)) if aux else None
        self.fc = vision.TcLinear(2048, num_classes) if f else nn.Linear(2048, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.mixed(self.stem(x))
        out = self.fc(torch.flatten(F.adaptive_avg_pool2d(self.tail(x), 1), 1))
        if self.aux is not None and self.training:
            a = self.aux(F.adaptive_avg_pool2d(x, 5))
            out = out + 0.4 * self.aux_fc(torch.flatten(F.adaptive_avg_pool2d(a, 1), 1))
        return out


# ---- MobileNet v1 ----------------------------------------------------------------------------------------------------
class MobileNetV1(nn.Module):
    """Depthwise-separable stack, width multiplier 0.25 (~0.47 M parameters; the table's 15 MB is the TF
    checkpoint with its optimizer slots). Depthwise convolutions are grouped cuDNN convolutions."""

    def __init__(self, width: float = 0.25, num_classes: int = 1000, fused: bool = True):
        super().__init__()
        def c(v: int) -> int:
            return max(8, int(v * width) // 8 * 8)
        cfg = [(64, 1), (128, 2), (128, 1), (256, 2), (256, 1), (512, 2), (512, 1), (512, 1), (512, 1), (512, 1),
               (512, 1), (1024, 2), (1024, 1)]
        layers: List[nn.Module] = [ConvBN(3, c(32), 3, 2, 1, fused=fused)]
        cin = c(32)
        for cout, stride in cfg:
            layers.append(ConvBN(cin, cin, 3, stride, 1, groups=cin, fused=fused))       # depthwise
            layers.append(ConvBN(cin, c(cout), 1, fused=fused))                          # pointwise
            cin = c(cout)
        self.features = nn.Sequential(*layers)
        self.fc = vision.TcLinear(cin, num_classes) if fused else nn.Linear(cin, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.fc(torch.flatten(F.adaptive_avg_pool2d(self.features(x), 1), 1))


def googlenet(num_classes: int = 1000, fused: bool = True) -> GoogLeNet:
    return GoogLeNet(num_classes, fused)


def inception3(num_classes: int = 1000, fused: bool = True) -> InceptionV3:
    return InceptionV3(num_classes, fused)


def mobilenet_v1_025(num_classes: int = 1000, fused: bool = True) -> MobileNetV1:
    return MobileNetV1(0.25, num_classes, fused)
