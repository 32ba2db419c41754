"""CNN job models named by the reference's model table (model/model_factory.py:4-55,
core/models.py:20): ResNet-50/101/152, VGG-11/16/19, AlexNet. Random-init, NHWC (channels_last),
bf16. Convolutions go through cuDNN; every fully connected layer runs on the tcgen05 GEMM
(:class:`tiresias_b200.ops.linear.TcLinear`) and every BatchNorm (+ residual add + ReLU) on the fused
kernels of :class:`tiresias_b200.ops.batchnorm.FusedBatchNormAct`. ``fused=False`` builds the same
network from stock PyTorch layers only (nn.BatchNorm2d, nn.Linear): the PyTorch+NCCL baseline arm.
"""
from __future__ import annotations

from typing import List, Sequence

import torch
from torch import nn

from ..ops import native
from ..ops.batchnorm import FusedBatchNormAct
from ..ops.conv import GradBox, TcConv1x1, TcConv2d, TcConv3x3Bias, _parse_ops, packed_stem_conv, park_skip_grad
from ..ops.linear import TcLinear
from ..ops.pooling import MaxPool, MaxPool3x3S2

import os

# Convolutions on our own tcgen05 kernels (ops/conv.py): 1x1 / stride 1 as plain GEMMs over the NHWC pixel
# rows, everything else with a multiple of 64 input channels as implicit GEMMs fed by im2col-mode TMA loads.
# Correct (tests/test_gpu_conv.py, test_gpu_trainer.py) but cuDNN wins every pass in situ on B200 at batch
# 256 (profiles/conv_r1g.md: +1 % of the ResNet-50 step per pointwise pass, +3-5 % per k x k pass), so the
# routing is opt-in:  TB200_TC_CONV="p:fdw,k:fdw"  (p = pointwise, k = the rest; f / d / w = forward, input
# gradient, weight gradient; "1" = everything; see ops.conv._parse_ops). TB200_TC_CONV1X1=1 is the older
# spelling of "p:fdw".
TC_CONV1X1 = os.environ.get("TB200_TC_CONV1X1", "0") == "1"
TC_CONV = _parse_ops(os.environ.get("TB200_TC_CONV", "1"))
if TC_CONV1X1:
    TC_CONV["p"] = set("fdw")


class TorchBNAct(nn.Module):
    """Stock-PyTorch counterpart of :class:`FusedBatchNormAct` (BatchNorm2d, then add, then ReLU as
    separate library kernels) — what the PyTorch+NCCL baseline arm runs."""

    def __init__(self, num_features: int, relu: bool = False):
        super().__init__()
        self.bn = nn.BatchNorm2d(num_features)
        self.relu = relu

    @property
    def weight(self) -> torch.Tensor:
        return self.bn.weight

    def forward(self, x: torch.Tensor, residual: torch.Tensor = None) -> torch.Tensor:
        y = self.bn(x)
        if residual is not None:
            y = y + residual
        return torch.relu_(y) if self.relu else y


def _bn(channels: int, relu: bool, fused: bool) -> nn.Module:
    return FusedBatchNormAct(channels, relu=relu) if fused else TorchBNAct(channels, relu=relu)


def _conv(cin: int, cout: int, k: int, stride: int = 1, pad: int = 0, fused: bool = False) -> nn.Module:
    """cuDNN convolution unless TB200_TC_CONV routes passes of this geometry to the tcgen05 kernels."""
    if fused and TC_CONV["p"] and k == 1 and stride == 1 and pad == 0 and cin % 8 == 0 and cout % 8 == 0:
        return TcConv1x1(cin, cout, TC_CONV["p"], emit_stats=True)
    if fused and TC_CONV["k"] and not (k == 1 and stride == 1) and cin % 64 == 0 and cout % 8 == 0:
        return TcConv2d(cin, cout, k, stride, pad, TC_CONV["k"], emit_stats=True)
    return nn.Conv2d(cin, cout, k, stride, pad, bias=False)


class Bottleneck(nn.Module):
    """ResNet v1.5 bottleneck (stride on the 3x3). Every BatchNorm carries its ReLU, and the last one
    also the residual add, so a block is 3-4 convolutions + 3-4 fused normalisation passes."""

    expansion = 4

    def __init__(self, cin: int, width: int, stride: int, fused: bool = True):
        super().__init__()
        cout = width * self.expansion
        self.conv1, self.bn1 = _conv(cin, width, 1, fused=fused), _bn(width, True, fused)
        self.conv2, self.bn2 = _conv(width, width, 3, stride, 1, fused=fused), _bn(width, True, fused)
        self.conv3, self.bn3 = _conv(width, cout, 1, fused=fused), _bn(cout, True, fused)
        self.down_conv = self.down_bn = None
        if stride != 1 or cin != cout:
            self.down_conv, self.down_bn = _conv(cin, cout, 1, stride, fused=fused), _bn(cout, False, fused)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if isinstance(self.conv1, TcConv1x1) and self.conv1.fuses_skip_grad(x):
            # Training on the tcgen05 path: the skip branch is built last, so its backward runs first and parks
            # its gradient w.r.t. x; conv1's dgrad GEMM adds it in the epilogue (ops/conv.py GradBox) instead of
            # autograd summing the two branches with an element-wise kernel.
            box = GradBox()
            out = self.bn1(self.conv1(x, grad_box=box))
            out = self.conv3(self.bn2(self.conv2(out)))
            idt = park_skip_grad(x, box)
            if self.down_conv is not None:
                idt = self.down_bn(self.down_conv(idt))
            return self.bn3(out, idt)
        idt = x if self.down_conv is None else self.down_bn(self.down_conv(x))
        out = self.bn1(self.conv1(x))
        out = self.bn2(self.conv2(out))
        return self.bn3(self.conv3(out), idt)


class ResNet(nn.Module):
    def __init__(self, depths: Sequence[int], num_classes: int = 1000, fused: bool = True):
        super().__init__()
        self.fused = fused
        self.stem_conv, self.stem_bn = _conv(3, 64, 7, 2, 3), _bn(64, True, fused)
        self.stem_pool = MaxPool3x3S2() if fused else nn.MaxPool2d(3, 2, 1)
        blocks: List[nn.Module] = []
        cin = 64
        for i, (n, width) in enumerate(zip(depths, (64, 128, 256, 512))):
            for j in range(n):
                blocks.append(Bottleneck(cin, width, 2 if (j == 0 and i > 0) else 1, fused))
                cin = width * Bottleneck.expansion
        self.blocks = nn.Sequential(*blocks)
        self.pool = nn.AdaptiveAvgPool2d(1)
        self.fc = TcLinear(cin, num_classes) if fused else nn.Linear(cin, num_classes)
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, TcConv1x1, TcConv2d)):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
        for m in self.modules():
            if isinstance(m, Bottleneck):
                nn.init.zeros_(m.bn3.weight)            # zero-init last BN of each block

    input_channels = 3          # the trainer may feed 8 (rgb + 5 zero channels) or the packed 64: see _stem

    @property
    def packed_stem(self) -> bool:
        """True when the trainer should feed ``native.stem7x7_pack`` output (the stem then runs on the tcgen05
        implicit GEMM, csrc/stem.cu) instead of the normalised image."""
        return self.fused and bool(TC_CONV["k"] >= {"f", "w"})

    def _stem(self, x: torch.Tensor) -> torch.Tensor:
        """7x7/2 convolution. A 64-channel input is the packed tensor of csrc/stem.cu: 4x1 / stride-1 implicit
        GEMM on our kernels. An 8-channel input (rgb00000, written by tb200_u8_to_bf16_norm_pad8) is
        convolved with the 3-channel weight zero-padded to 8 input channels — same result, but cuDNN can
        use its 8-channel-aligned sm_100 kernels instead of the legacy 3-channel one."""
        conv = self.stem_conv
        if x.shape[1] == 64 and x.is_cuda:
            return packed_stem_conv(x, conv.weight, emit_stats=self.training)
        if x.shape[1] == 8:
            w8 = torch.nn.functional.pad(conv.weight, (0, 0, 0, 0, 0, 5))
            return torch.nn.functional.conv2d(x, w8, None, conv.stride, conv.padding)
        return conv(x)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.stem_pool(self.stem_bn(self._stem(x)))
        x = self.blocks(x)
        return self.fc(torch.flatten(self.pool(x), 1))


def resnet50(num_classes: int = 1000, fused: bool = True) -> ResNet:
    return ResNet((3, 4, 6, 3), num_classes, fused)


def resnet101(num_classes: int = 1000, fused: bool = True) -> ResNet:
    return ResNet((3, 4, 23, 3), num_classes, fused)


def resnet152(num_classes: int = 1000, fused: bool = True) -> ResNet:
    return ResNet((3, 8, 36, 3), num_classes, fused)


class _TorchLinearAct(nn.Module):
    def __init__(self, cin: int, cout: int, relu: bool):
        super().__init__()
        self.fc, self.relu = nn.Linear(cin, cout), relu

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = self.fc(x)
        return torch.relu_(y) if self.relu else y


def _classifier(cin: int, num_classes: int, fused: bool):
    """fc6 / fc7 / fc8 of VGG and AlexNet: tcgen05 GEMMs with the bias+ReLU epilogue, or stock
    nn.Linear + ReLU for the baseline arm."""
    if fused:
        return (TcLinear(cin, 4096, act=native.ACT_RELU), TcLinear(4096, 4096, act=native.ACT_RELU),
                TcLinear(4096, num_classes))
    return (_TorchLinearAct(cin, 4096, True), _TorchLinearAct(4096, 4096, True),
            _TorchLinearAct(4096, num_classes, False))


_VGG_CFG = {
    11: (64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"),
    16: (64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"),
    19: (64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M",
         512, 512, 512, 512, "M"),
}


class VGG(nn.Module):
    """fc6 (25088x4096) is the 392 MB tensor of core/models.py:9 - the bandwidth-bound bucket."""

    def __init__(self, depth: int, num_classes: int = 1000, fused: bool = True):
        super().__init__()
        layers: List[nn.Module] = []
        cin = 3
        for v in _VGG_CFG[depth]:
            if v == "M":
                layers.append(MaxPool(2, 2) if fused else nn.MaxPool2d(2, 2))
            elif fused and cin % 64 == 0 and v % 64 == 0:
                layers.append(TcConv3x3Bias(cin, v, relu=True))      # bias + ReLU in the convolution's epilogue
                cin = v
            else:
                layers += [nn.Conv2d(cin, v, 3, 1, 1), nn.ReLU(inplace=True)]
                cin = v
        self.features = nn.Sequential(*layers)
        self.fc6, self.fc7, self.fc8 = _classifier(512 * 7 * 7, num_classes, fused)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.features(x)
        # NHWC memory order -> flatten in (h, w, c) order so the view is free
        x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
        return self.fc8(self.fc7(self.fc6(x)))


def vgg11(num_classes: int = 1000, fused: bool = True) -> VGG:
    return VGG(11, num_classes, fused)


def vgg16(num_classes: int = 1000, fused: bool = True) -> VGG:
    return VGG(16, num_classes, fused)


def vgg19(num_classes: int = 1000, fused: bool = True) -> VGG:
    return VGG(19, num_classes, fused)


class AlexNet(nn.Module):
    """fc layers 144/64/15.6 MB per core/models.py:11 (9216x4096, 4096x4096, 4096x1000)."""

    def __init__(self, num_classes: int = 1000, fused: bool = True):
        super().__init__()
        pool = (lambda: MaxPool(3, 2)) if fused else (lambda: nn.MaxPool2d(3, 2))

        def c3(cin, cout):
            if fused:
                return (TcConv3x3Bias(cin, cout, relu=True),)
            return (nn.Conv2d(cin, cout, 3, 1, 1), nn.ReLU(inplace=True))

        self.features = nn.Sequential(
            nn.Conv2d(3, 64, 11, 4, 2), nn.ReLU(inplace=True), pool(),
            nn.Conv2d(64, 192, 5, 1, 2), nn.ReLU(inplace=True), pool(),
            *c3(192, 384), *c3(384, 256), *c3(256, 256), pool())
        self.fc6, self.fc7, self.fc8 = _classifier(256 * 6 * 6, num_classes, fused)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.features(x)
        x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
        return self.fc8(self.fc7(self.fc6(x)))


def alexnet(num_classes: int = 1000, fused: bool = True) -> AlexNet:
    return AlexNet(num_classes, fused)
