"""MNIST-shaped CNN: conv5x5(1->32) ReLU pool2 conv5x5(32->64) ReLU pool2
fc(3136->2048) ReLU fc(2048->10).

Same architecture and parameter names as the reference example
(p2pfl/learning/pytorch/mnist_examples/models/cnn.py:55-91): 8 tensors,
6 497 162 parameters, 25 988 648 bytes in fp32.
"""

from __future__ import annotations

from typing import Optional

import torch
from torch import nn


class CNN(nn.Module):
    lr = 1e-3
    optimizer = "adam"
    in_shape = (1, 28, 28)
    num_classes = 10
    fused_spec = "cnn"

    def __init__(self, in_channels: int = 1, out_channels: int = 10, seed: Optional[int] = None):
        super().__init__()
        if seed is not None:
            torch.manual_seed(seed)
        self.conv1 = nn.Conv2d(in_channels, 32, kernel_size=5, padding="same")
        self.conv2 = nn.Conv2d(32, 64, kernel_size=5, padding="same")
        self.l1 = nn.Linear(7 * 7 * 64, 2048)
        self.l2 = nn.Linear(2048, out_channels)
        self.num_classes = out_channels

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = torch.max_pool2d(torch.relu(self.conv1(x)), 2, 2)
        x = torch.max_pool2d(torch.relu(self.conv2(x)), 2, 2)
        x = x.reshape(x.shape[0], -1)
        x = torch.relu(self.l1(x))
        return self.l2(x)
