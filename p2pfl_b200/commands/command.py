"""Command pattern for inbound messages (reference: p2pfl/commands/command.py:24-43)."""

from __future__ import annotations

import abc
from typing import Any


class Command(abc.ABC):
    """A named handler; the protocol engine dispatches on :meth:`get_name`."""

    @staticmethod
    @abc.abstractmethod
    def get_name() -> str: ...

    @abc.abstractmethod
    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None: ...
