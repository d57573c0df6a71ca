"""ResNet-18 / 34 / 50 / 101 for the BASELINE.json configs (ResNet-18 CIFAR-shaped,
ResNet-50 ImageNet-shaped).

The reference's own CNN (``models.sr_4k_2x.Net``, Fairscale-DDP.py:13,74) is absent;
BASELINE.json substitutes ResNet-18/50 on synthetic data.  Parameter names, shapes
and counts match torchvision's (``resnet18``: 11,689,512; ``resnet50``: 25,557,032;
SURVEY §2.2 E23) so checkpoints interchange.

B200-first: every conv → norm → (residual add) → ReLU chain goes through
``ops.layers``: on CUDA in channels-last half precision the norm + add + activation is
one fused kernel reading the conv output once (W8), the stride-1 1×1 convolutions are
plain GEMMs on the tcgen05 kernel (W7), and the classifier is the fused linear.  On CPU
the same modules fall back to stock torch ops, which is what the gloo tests exercise.
"""
from __future__ import annotations

from typing import List, Optional, Type, Union

import torch
import torch.nn as nn

from ..ops.layers import BatchNormAct2d, Conv2d, Linear, MaxPool2d

__all__ = ["ResNet", "resnet18", "resnet34", "resnet50", "resnet101"]


def conv3x3(inp: int, out: int, stride: int = 1) -> Conv2d:
    return Conv2d(inp, out, kernel_size=3, stride=stride, padding=1, bias=False)


def conv1x1(inp: int, out: int, stride: int = 1) -> Conv2d:
    return Conv2d(inp, out, kernel_size=1, stride=stride, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None):
        super().__init__()
        self.conv1 = conv3x3(inplanes, planes, stride)
        self.bn1 = BatchNormAct2d(planes, act="relu")
        self.conv2 = conv3x3(planes, planes)
        self.bn2 = BatchNormAct2d(planes, act="relu")  # ReLU applied after the residual add
        self.downsample = downsample

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x))
        return self.bn2(self.conv2(out), residual=identity)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None):
        super().__init__()
        self.conv1 = conv1x1(inplanes, planes)
        self.bn1 = BatchNormAct2d(planes, act="relu")
        self.conv2 = conv3x3(planes, planes, stride)
        self.bn2 = BatchNormAct2d(planes, act="relu")
        self.conv3 = conv1x1(planes, planes * self.expansion)
        self.bn3 = BatchNormAct2d(planes * self.expansion, act="relu")
        self.downsample = downsample

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        # conv1 hands the block input back for the bypass branch (ops.layers.Conv2d.forward_carry)
        out, x = self.conv1.forward_carry(x)
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn2(self.conv2(self.bn1(out)))
        return self.bn3(self.conv3(out), residual=identity)


class ResNet(nn.Module):
    def __init__(self, block: Type[Union[BasicBlock, Bottleneck]], layers: List[int], num_classes: int = 1000,
                 small_input: bool = False, zero_init_residual: bool = False):
        """``small_input``: CIFAR-style stem (3×3 stride-1 conv, no max-pool) for 32×32 images."""
        super().__init__()
        self.inplanes = 64
        if small_input:
            self.conv1 = Conv2d(3, 64, kernel_size=3, stride=1, padding=1, bias=False)
            self.maxpool: nn.Module = nn.Identity()
        else:
            self.conv1 = Conv2d(3, 64, kernel_size=7, stride=2, padding=3, bias=False)
            self.maxpool = MaxPool2d(kernel_size=3, stride=2, padding=1)
        self.bn1 = BatchNormAct2d(64, act="relu")
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = Linear(512 * block.expansion, num_classes)

        from ..ops.layers import pair_conv_bn
        pair_conv_bn(self)          # conv epilogues emit the statistics of the BatchNorm that follows
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)
        if zero_init_residual:
            for m in self.modules():
                if isinstance(m, Bottleneck):
                    nn.init.constant_(m.bn3.weight, 0)
                elif isinstance(m, BasicBlock):
                    nn.init.constant_(m.bn2.weight, 0)

    def _make_layer(self, block, planes: int, blocks: int, stride: int = 1) -> nn.Sequential:
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(conv1x1(self.inplanes, planes * block.expansion, stride),
                                       BatchNormAct2d(planes * block.expansion, act=None))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes))
        return nn.Sequential(*layers)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.maxpool(self.bn1(self.conv1(x)))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = torch.flatten(self.avgpool(x), 1)
        return self.fc(x)


def resnet18(num_classes: int = 1000, **kw) -> ResNet:
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes, **kw)


def resnet34(num_classes: int = 1000, **kw) -> ResNet:
    return ResNet(BasicBlock, [3, 4, 6, 3], num_classes, **kw)


def resnet50(num_classes: int = 1000, **kw) -> ResNet:
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes, **kw)


def resnet101(num_classes: int = 1000, **kw) -> ResNet:
    return ResNet(Bottleneck, [3, 4, 23, 3], num_classes, **kw)
