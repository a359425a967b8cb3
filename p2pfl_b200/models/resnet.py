"""CIFAR-10-shaped ResNet-18 (BASELINE.json config 3).

The torchvision ResNet-18 layout with the CIFAR stem (3x3 conv, no max-pool):
122 ``state_dict`` entries, 11.17 M float parameters plus 20 int64
``num_batches_tracked`` counters which stay out of the arena.
Not part of the reference; required by BASELINE.json.
"""

from __future__ import annotations

from typing import Optional

import torch
from torch import nn


class BasicBlock(nn.Module):
    def __init__(self, cin: int, cout: int, stride: int) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.downsample: Optional[nn.Sequential] = None
        if stride != 1 or cin != cout:
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        idt = x if self.downsample is None else self.downsample(x)
        out = torch.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return torch.relu(out + idt)


class ResNet18(nn.Module):
    lr = 1e-3
    optimizer = "adam"
    in_shape = (3, 32, 32)
    num_classes = 10
    fused_spec = "resnet18"

    def __init__(self, num_classes: int = 10, seed: Optional[int] = None) -> None:
        super().__init__()
        if seed is not None:
            torch.manual_seed(seed)
        self.conv1 = nn.Conv2d(3, 64, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        cfg = [(64, 64, 1), (64, 64, 1), (64, 128, 2), (128, 128, 1), (128, 256, 2), (256, 256, 1), (256, 512, 2), (512, 512, 1)]
        self.layer1 = nn.Sequential(BasicBlock(*cfg[0]), BasicBlock(*cfg[1]))
        self.layer2 = nn.Sequential(BasicBlock(*cfg[2]), BasicBlock(*cfg[3]))
        self.layer3 = nn.Sequential(BasicBlock(*cfg[4]), BasicBlock(*cfg[5]))
        self.layer4 = nn.Sequential(BasicBlock(*cfg[6]), BasicBlock(*cfg[7]))
        self.fc = nn.Linear(512, num_classes)
        self.num_classes = num_classes

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = torch.relu(self.bn1(self.conv1(x)))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = x.mean(dim=(2, 3))
        return self.fc(x)
