"""ResNet-18 / ResNet-50 (north-star configs: CIFAR-10-shaped 32x32 and ImageNet-shaped 224x224;
``BASELINE.json:configs[2,3]`` - not part of the reference, which only has the ConvNet).

Module/parameter names follow the usual torchvision layout (``conv1, bn1, layer1..4.{i}.conv{j}/bn{j},
downsample.0/1, fc``) so a state_dict interchanges with the library model the baseline harness runs.
Convolutions are bias-free and channels-last; BatchNorm+ReLU is one fused kernel; the residual
add+ReLU is fused; ``sync_bn=True`` swaps in SyncBatchNorm (statistics reduced over NVLink P2P)."""
from __future__ import annotations

from typing import List, Optional, Type

import torch
import torch.nn.functional as F
from torch import nn

from .. import ops


def _bn(c: int, relu: bool, sync: bool) -> nn.Module:
    cls = ops.SyncBatchNorm if sync else ops.BatchNorm2d
    return cls(c, fuse_relu=relu)


def _add_relu(x: torch.Tensor, res: torch.Tensor) -> torch.Tensor:
    if x.is_cuda:
        from ..ops import cuda
        return cuda.add_relu(x, res)
    return F.relu(x + res)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None, sync_bn: bool = False):
        super().__init__()
        self.conv1 = ops.Conv2d(cin, planes, 3, stride, 1, bias=False)
        self.bn1 = _bn(planes, True, sync_bn)
        self.conv2 = ops.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = _bn(planes, False, sync_bn)
        self.downsample = downsample

    def forward(self, x):
        idt = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x))
        out = self.bn2(self.conv2(out))
        return _add_relu(out, idt)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None, sync_bn: bool = False):
        super().__init__()
        self.conv1 = ops.Conv2d(cin, planes, 1, 1, 0, bias=False)
        self.bn1 = _bn(planes, True, sync_bn)
        self.conv2 = ops.Conv2d(planes, planes, 3, stride, 1, bias=False)
        self.bn2 = _bn(planes, True, sync_bn)
        self.conv3 = ops.Conv2d(planes, planes * 4, 1, 1, 0, bias=False)
        self.bn3 = _bn(planes * 4, False, sync_bn)
        self.downsample = downsample

    def forward(self, x):
        idt = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x))
        out = self.bn2(self.conv2(out))
        out = self.bn3(self.conv3(out))
        return _add_relu(out, idt)


class ResNet(nn.Module):
    def __init__(self, block: Type[nn.Module], layers: List[int], num_classes: int = 1000, sync_bn: bool = False):
        super().__init__()
        self.inplanes, self.sync_bn = 64, sync_bn
        self.compute_dtype = None   # torch.bfloat16: bf16 activations / convs with fp32 master weights and fp32 BN parameters
        self.conv1 = ops.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = _bn(64, True, sync_bn)
        self.maxpool = ops.MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(block, 64, layers[0], 1)
        self.layer2 = self._make_layer(block, 128, layers[1], 2)
        self.layer3 = self._make_layer(block, 256, layers[2], 2)
        self.layer4 = self._make_layer(block, 512, layers[3], 2)
        self.fc = ops.Linear(512 * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def _make_layer(self, block, planes: int, n: int, stride: int) -> nn.Sequential:
        down = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            down = nn.Sequential(ops.Conv2d(self.inplanes, planes * block.expansion, 1, stride, 0, bias=False),
                                 _bn(planes * block.expansion, False, self.sync_bn))
        blocks = [block(self.inplanes, planes, stride, down, self.sync_bn)]
        self.inplanes = planes * block.expansion
        blocks += [block(self.inplanes, planes, sync_bn=self.sync_bn) for _ in range(1, n)]
        return nn.Sequential(*blocks)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.compute_dtype is not None:
            x = x.to(self.compute_dtype)
        x = x.contiguous(memory_format=torch.channels_last)
        x = self.maxpool(self.bn1(self.conv1(x)))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        if x.is_cuda:
            from ..ops import cuda
            x = cuda.global_avg_pool(x)
        else:
            x = x.mean((2, 3))
        return self.fc(x)


def resnet18(num_classes: int = 10, sync_bn: bool = False) -> ResNet:
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes, sync_bn)


def resnet50(num_classes: int = 1000, sync_bn: bool = False) -> ResNet:
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes, sync_bn)
