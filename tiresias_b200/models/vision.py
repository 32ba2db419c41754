"""CNN job models named by the reference's model table (model/model_factory.py:4-55,
core/models.py:20): ResNet-50/101/152, VGG-11/16/19, AlexNet. Random-init, NHWC (channels_last),
bf16. Convolutions go through cuDNN; every fully connected layer runs on the tcgen05 GEMM
(:class:`tiresias_b200.ops.linear.TcLinear`).
"""
from __future__ import annotations

from typing import List, Sequence

import torch
from torch import nn

from ..ops import native
from ..ops.linear import TcLinear


def _conv_bn(cin: int, cout: int, k: int, stride: int = 1, pad: int = 0) -> List[nn.Module]:
    return [nn.Conv2d(cin, cout, k, stride, pad, bias=False), nn.BatchNorm2d(cout)]


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin: int, width: int, stride: int):
        super().__init__()
        cout = width * self.expansion
        self.body = nn.Sequential(
            *_conv_bn(cin, width, 1), nn.ReLU(inplace=True),
            *_conv_bn(width, width, 3, stride, 1), nn.ReLU(inplace=True),     # v1.5: stride on 3x3
            *_conv_bn(width, cout, 1))
        self.down = None
        if stride != 1 or cin != cout:
            self.down = nn.Sequential(*_conv_bn(cin, cout, 1, stride))
        self.relu = nn.ReLU(inplace=True)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        idt = x if self.down is None else self.down(x)
        return self.relu(self.body(x) + idt)


class ResNet(nn.Module):
    def __init__(self, depths: Sequence[int], num_classes: int = 1000):
        super().__init__()
        self.stem = nn.Sequential(*_conv_bn(3, 64, 7, 2, 3), nn.ReLU(inplace=True),
                                  nn.MaxPool2d(3, 2, 1))
        blocks: List[nn.Module] = []
        cin = 64
        for i, (n, width) in enumerate(zip(depths, (64, 128, 256, 512))):
            for j in range(n):
                blocks.append(Bottleneck(cin, width, 2 if (j == 0 and i > 0) else 1))
                cin = width * Bottleneck.expansion
        self.blocks = nn.Sequential(*blocks)
        self.pool = nn.AdaptiveAvgPool2d(1)
        self.fc = TcLinear(cin, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
        for m in self.modules():
            if isinstance(m, Bottleneck):
                nn.init.zeros_(m.body[-1].weight)       # zero-init last BN of each block

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.blocks(self.stem(x))
        return self.fc(torch.flatten(self.pool(x), 1))


def resnet50(num_classes: int = 1000) -> ResNet:
    return ResNet((3, 4, 6, 3), num_classes)


def resnet101(num_classes: int = 1000) -> ResNet:
    return ResNet((3, 4, 23, 3), num_classes)


def resnet152(num_classes: int = 1000) -> ResNet:
    return ResNet((3, 8, 36, 3), num_classes)


_VGG_CFG = {
    11: (64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"),
    16: (64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"),
    19: (64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M",
         512, 512, 512, 512, "M"),
}


class VGG(nn.Module):
    """fc6 (25088x4096) is the 392 MB tensor of core/models.py:9 - the bandwidth-bound bucket."""

    def __init__(self, depth: int, num_classes: int = 1000):
        super().__init__()
        layers: List[nn.Module] = []
        cin = 3
        for v in _VGG_CFG[depth]:
            if v == "M":
                layers.append(nn.MaxPool2d(2, 2))
            else:
                layers += [nn.Conv2d(cin, v, 3, 1, 1), nn.ReLU(inplace=True)]
                cin = v
        self.features = nn.Sequential(*layers)
        self.fc6 = TcLinear(512 * 7 * 7, 4096, act=native.ACT_RELU)
        self.fc7 = TcLinear(4096, 4096, act=native.ACT_RELU)
        self.fc8 = TcLinear(4096, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.features(x)
        # NHWC memory order -> flatten in (h, w, c) order so the view is free
        x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
        return self.fc8(self.fc7(self.fc6(x)))


def vgg11(num_classes: int = 1000) -> VGG:
    return VGG(11, num_classes)


def vgg16(num_classes: int = 1000) -> VGG:
    return VGG(16, num_classes)


def vgg19(num_classes: int = 1000) -> VGG:
    return VGG(19, num_classes)


class AlexNet(nn.Module):
    """fc layers 144/64/15.6 MB per core/models.py:11 (9216x4096, 4096x4096, 4096x1000)."""

    def __init__(self, num_classes: int = 1000):
        super().__init__()
        self.features = nn.Sequential(
            nn.Conv2d(3, 64, 11, 4, 2), nn.ReLU(inplace=True), nn.MaxPool2d(3, 2),
            nn.Conv2d(64, 192, 5, 1, 2), nn.ReLU(inplace=True), nn.MaxPool2d(3, 2),
            nn.Conv2d(192, 384, 3, 1, 1), nn.ReLU(inplace=True),
            nn.Conv2d(384, 256, 3, 1, 1), nn.ReLU(inplace=True),
            nn.Conv2d(256, 256, 3, 1, 1), nn.ReLU(inplace=True), nn.MaxPool2d(3, 2))
        self.fc6 = TcLinear(256 * 6 * 6, 4096, act=native.ACT_RELU)
        self.fc7 = TcLinear(4096, 4096, act=native.ACT_RELU)
        self.fc8 = TcLinear(4096, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.features(x)
        x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
        return self.fc8(self.fc7(self.fc6(x)))


def alexnet(num_classes: int = 1000) -> AlexNet:
    return AlexNet(num_classes)
