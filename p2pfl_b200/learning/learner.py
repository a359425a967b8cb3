"""Learner template (reference: p2pfl/learning/learner.py:24-150).

The stage machine only talks to this interface.  ``check_parameters`` and
``export_model`` are additions: the first validates a decoded model against
the local layout before it reaches the aggregator (Q20), the second returns
the cheapest representation the learner has (a flat arena view on the device
path, a state-dict otherwise).
"""

from __future__ import annotations

import abc
from typing import Any, Dict, Optional, Tuple


class NodeLearner(abc.ABC):
    """Strategy interface for local training."""

    @abc.abstractmethod
    def __init__(self, model: Any, data: Any, self_addr: str, epochs: int) -> None: ...

    @abc.abstractmethod
    def set_model(self, model: Any) -> None: ...

    @abc.abstractmethod
    def set_data(self, data: Any) -> None: ...

    @abc.abstractmethod
    def encode_parameters(self, params: Any = None) -> Any:
        """Serialise ``params`` (default: the live model) for the wire."""

    @abc.abstractmethod
    def decode_parameters(self, data: Any) -> Any:
        """Inverse of :meth:`encode_parameters`; raises ``DecodingParamsError``."""

    @abc.abstractmethod
    def check_parameters(self, params: Any) -> None:
        """Raise ``ModelNotMatchingError`` unless ``params`` fits the local model."""

    @abc.abstractmethod
    def set_parameters(self, params: Any) -> None: ...

    @abc.abstractmethod
    def get_parameters(self) -> Dict[str, Any]:
        """``{name: tensor}`` in ``state_dict`` order."""

    def export_model(self) -> Any:
        """Model object to hand to the aggregator (default: state-dict)."""
        return self.get_parameters()

    @abc.abstractmethod
    def set_epochs(self, epochs: int) -> None: ...

    @abc.abstractmethod
    def fit(self) -> None: ...

    @abc.abstractmethod
    def interrupt_fit(self) -> None: ...

    @abc.abstractmethod
    def evaluate(self) -> Dict[str, float]: ...

    @abc.abstractmethod
    def get_num_samples(self) -> Tuple[int, int]:
        """``(train samples, test samples)``."""

    def close(self) -> None:
        """Release device resources (optional)."""

    def begin_round(self, round: Optional[int]) -> None:
        """Hook called by TrainStage before evaluate/fit (optional)."""
