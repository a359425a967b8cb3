"""Stage base and run context (reference: p2pfl/stages/stage.py:23-34).

The reference passes a bag of ``**kwargs`` to every stage
(p2pfl/node.py:349-359); here the same items travel in one typed
``LearningContext``.
"""

from __future__ import annotations

import abc
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, Optional, Type


@dataclass
class LearningContext:
    rounds: int
    epochs: int
    state: Any  # NodeState
    model: Any
    data: Any
    communication_protocol: Any
    aggregator: Any
    learner_class: Any
    early_stopping_fn: Callable[[], bool]
    learner_kwargs: Dict[str, Any] = field(default_factory=dict)
    timings: Dict[str, float] = field(default_factory=dict)  # per-stage host seconds


class Stage(abc.ABC):
    """A step of the learning workflow; ``execute`` returns the next stage class
    or ``None`` to end the workflow."""

    @staticmethod
    @abc.abstractmethod
    def name() -> str: ...

    @staticmethod
    @abc.abstractmethod
    def execute(ctx: LearningContext) -> Optional[Type["Stage"]]: ...
