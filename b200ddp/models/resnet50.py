"""ResNet-50 (BASELINE.json config 3: "ResNet-50 DDP on synthetic 224x224 batches").

Not part of the reference repository; the comparator is torchvision's ``resnet50``
under the reference's DDP loop.  Same architecture (v1.5: stride on
the 3x3 convolution), same 25,557,032 parameters and the same state-dict keys as
torchvision, so weights are interchangeable.

What the framework contributes for this config: the flat parameter / gradient /
buffer arenas, bucketed overlap with backward, the fused NVLS all-reduce +
optimizer kernel per bucket, one coalesced buffer broadcast per step for the
BatchNorm running statistics, bf16 compute with fp32 master
weights, the fused cross-entropy kernel, and every convolution (forward, dgrad,
wgrad) and the classifier on the tcgen05 GEMM: 1x1 convolutions feed the NHWC
activation to the GEMM directly, 3x3/7x7 and strided ones go through the
im2col/col2im gather kernels (``csrc/conv.cu``); BatchNorm is our kernel too, fused
with ReLU and the residual add (``csrc/batchnorm.cu``: 2 passes forward, 2 backward).
Max and global average pooling are NHWC gather kernels too (csrc/pool.cu); torch's on the CPU.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class _BN(nn.BatchNorm2d):
    """BatchNorm with fp32 statistics/affine math on bf16 activations (cuDNN mixed mode);
    the affine parameters may live in a bf16 arena."""

    def forward(self, x, relu: bool = False, residual=None):
        from ..ops import functional as Fn
        if self.training and self.track_running_stats and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
        return Fn.batch_norm_act(x, self.weight, self.bias, self.running_mean, self.running_var,
                                 self.training, self.momentum, self.eps, relu=relu, residual=residual)


class _Conv(nn.Conv2d):
    """nn.Conv2d container (same state-dict keys as torchvision) whose forward runs on the
    tcgen05 GEMM through im2col/col2im (ops/functional.py) on a B200, cuDNN elsewhere."""

    def forward(self, x):
        from ..ops import functional as Fn
        return Fn.conv2d(x, self.weight, self.stride[0], self.padding[0])


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: nn.Module | None = None):
        super().__init__()
        self.conv1 = _Conv(inplanes, planes, 1, bias=False)
        self.bn1 = _BN(planes)
        self.conv2 = _Conv(planes, planes, 3, stride=stride, padding=1, bias=False)
        self.bn2 = _BN(planes)
        self.conv3 = _Conv(planes, planes * 4, 1, bias=False)
        self.bn3 = _BN(planes * 4)
        self.downsample = downsample

    def forward(self, x):
        idt = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x), relu=True)
        out = self.bn2(self.conv2(out), relu=True)
        return self.bn3(self.conv3(out), relu=True, residual=idt)  # add + ReLU fused into the BN pass


class ResNet50(nn.Module):
    def __init__(self, num_classes: int = 1000):
        super().__init__()
        self.inplanes = 64
        self.conv1 = _Conv(3, 64, 7, stride=2, padding=3, bias=False)
        self.bn1 = _BN(64)
        self.maxpool = nn.MaxPool2d(3, stride=2, padding=1)
        self.layer1 = self._make_layer(64, 3)
        self.layer2 = self._make_layer(128, 4, stride=2)
        self.layer3 = self._make_layer(256, 6, stride=2)
        self.layer4 = self._make_layer(512, 3, stride=2)
        self.fc = nn.Linear(2048, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def _make_layer(self, planes: int, blocks: int, stride: int = 1) -> nn.Sequential:
        down = None
        if stride != 1 or self.inplanes != planes * 4:
            down = nn.Sequential(_Conv(self.inplanes, planes * 4, 1, stride=stride, bias=False),
                                 _BN(planes * 4))
        layers = [Bottleneck(self.inplanes, planes, stride, down)]
        self.inplanes = planes * 4
        layers += [Bottleneck(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    def forward(self, x):
        from ..ops import functional as Fn
        w = self.conv1.weight
        if x.dtype != w.dtype:
            x = x.to(w.dtype)
        if x.is_cuda:
            x = x.contiguous(memory_format=torch.channels_last)
        x = Fn.max_pool2d(self.bn1(self.conv1(x), relu=True), 3, 2, 1)
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = Fn.global_avg_pool(x)
        return Fn.linear(x, self.fc.weight, self.fc.bias)


def resnet50(num_classes: int = 1000) -> ResNet50:
    return ResNet50(num_classes)
