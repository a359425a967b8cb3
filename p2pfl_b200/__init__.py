"""p2pfl_b200: a B200-native decentralized federated-learning framework.

Public API mirrors the reference p2pfl package (Node, Settings, utils, logger)
- see DESIGN.md for the architecture and SURVEY.md for the parity map.
"""

__version__ = "0.1.0"

from p2pfl_b200.settings import Settings  # noqa: F401

__all__ = ["Settings", "__version__"]
