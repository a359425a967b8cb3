from p2pfl_b200.learning.data.synthetic import (
    Cifar10FederatedDM,
    MnistFederatedDM,
    SyntheticFederatedDM,
    TokenFederatedDM,
)

__all__ = ["SyntheticFederatedDM", "MnistFederatedDM", "Cifar10FederatedDM", "TokenFederatedDM"]
