"""Stand-in for the two ``torchmetrics`` names the reference models import
(mlp.py:26, cnn.py:26): ``Metric`` and multiclass ``Accuracy``."""

import torch
from torch import nn


class Metric(nn.Module):
    pass


class Accuracy(Metric):
    def __init__(self, task: str = "multiclass", num_classes: int = 10, **kw) -> None:
        super().__init__()
        self.num_classes = num_classes

    def forward(self, preds: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        if preds.dim() > target.dim():
            preds = preds.argmax(dim=1)
        return (preds == target).float().mean()
