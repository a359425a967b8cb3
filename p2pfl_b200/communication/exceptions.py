"""Communication exceptions (reference: p2pfl/communication/exceptions.py)."""

from p2pfl_b200.exceptions import CommunicationError, NeighborNotConnectedError

__all__ = ["CommunicationError", "NeighborNotConnectedError"]
