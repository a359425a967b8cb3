"""Model zoo: the reference's two example models plus the two BASELINE.json adds."""

from p2pfl_b200.models.cnn import CNN
from p2pfl_b200.models.gpt2 import GPT2, gpt2_small, gpt2_tiny
from p2pfl_b200.models.mlp import MLP
from p2pfl_b200.models.resnet import ResNet18

__all__ = ["MLP", "CNN", "ResNet18", "GPT2", "gpt2_small", "gpt2_tiny"]
