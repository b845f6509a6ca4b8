"""ResNet-18 for CIFAR-shaped (3x32x32) input [north star: BASELINE.json configs 2 and 4].

Parameter layout and names follow torchvision's ``resnet18(num_classes=10)`` (11 181 642
parameters) so the torchvision model is a drop-in oracle and checkpoints are
interchangeable.  BatchNorm running statistics are *buffers*; how federated aggregation
treats them is decided in ``fl/aggregators.py`` (mean of buffers, robust rule on params only).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin: int, cout: int, stride: int = 1):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.downsample = None
        if stride != 1 or cin != cout:
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        idt = x if self.downsample is None else self.downsample(x)
        y = F.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        return F.relu(y + idt)


class ResNet18(nn.Module):
    def __init__(self, num_classes: int = 10, in_ch: int = 3):
        super().__init__()
        self.conv1 = nn.Conv2d(in_ch, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        widths, cin = (64, 128, 256, 512), 64
        for i, w in enumerate(widths, start=1):
            stride = 1 if i == 1 else 2
            setattr(self, f"layer{i}", nn.Sequential(BasicBlock(cin, w, stride), BasicBlock(w, w, 1)))
            cin = w
        self.fc = nn.Linear(512, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.bn1(self.conv1(x))), 3, 2, 1)
        for i in range(1, 5):
            x = getattr(self, f"layer{i}")(x)
        return self.fc(torch.flatten(F.adaptive_avg_pool2d(x, 1), 1))


def resnet18(num_classes: int = 10) -> ResNet18:
    return ResNet18(num_classes)


def split_resnet18_two_stage(model: ResNet18) -> tuple[nn.Module, nn.Module]:
    """2-stage model-parallel split (BASELINE.json config 5): stem+layer1+layer2 | layer3+layer4+fc."""

    class Front(nn.Module):
        def __init__(s):
            super().__init__()
            s.conv1, s.bn1, s.layer1, s.layer2 = model.conv1, model.bn1, model.layer1, model.layer2

        def forward(s, x):
            x = F.max_pool2d(F.relu(s.bn1(s.conv1(x))), 3, 2, 1)
            return s.layer2(s.layer1(x))

    class Back(nn.Module):
        def __init__(s):
            super().__init__()
            s.layer3, s.layer4, s.fc = model.layer3, model.layer4, model.fc

        def forward(s, x):
            x = s.layer4(s.layer3(x))
            return s.fc(torch.flatten(F.adaptive_avg_pool2d(x, 1), 1))

    return Front(), Back()
