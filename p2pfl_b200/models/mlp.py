"""MNIST-shaped MLP: 784 -> 256 -> ReLU -> 128 -> ReLU -> 10.

Same architecture, parameter names (``l1``, ``l2``, ``l3``) and therefore the
same ``state_dict`` order/shapes as the reference example model
(p2pfl/learning/pytorch/mnist_examples/models/mlp.py:53-69), so host-path
payloads are interchangeable: 6 tensors, 235 146 parameters.  The reference
applies ``log_softmax`` before ``cross_entropy`` (idempotent); logits are
returned raw here and the loss kernel does the (single) log-softmax.
"""

from __future__ import annotations

from typing import Optional

import torch
from torch import nn


class MLP(nn.Module):
    lr = 1e-3
    optimizer = "adam"
    in_shape = (1, 28, 28)
    num_classes = 10
    # consumed by the fused B200 learner: (kind, name, in, out, relu)
    fused_spec = "mlp"

    def __init__(self, in_features: int = 784, hidden=(256, 128), out_channels: int = 10, seed: Optional[int] = None):
        super().__init__()
        if seed is not None:
            torch.manual_seed(seed)
        self.l1 = nn.Linear(in_features, hidden[0])
        self.l2 = nn.Linear(hidden[0], hidden[1])
        self.l3 = nn.Linear(hidden[1], out_channels)
        self.num_classes = out_channels

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x.reshape(x.shape[0], -1)
        x = torch.relu(self.l1(x))
        x = torch.relu(self.l2(x))
        return self.l3(x)
