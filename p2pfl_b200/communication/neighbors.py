"""Neighbour table.

One table type for every transport: the transport only supplies
``open_link`` / ``close_link`` callables.  Semantics follow the reference's
``Neighbors`` (p2pfl/communication/neighbors.py:27-170): direct neighbours own
a link, non-direct ones are names learned from flooded heartbeats; adding self
or a duplicate is refused; all accesses are taken under one lock (the
reference reads without it).
"""

from __future__ import annotations

import threading
import time
from dataclasses import dataclass
from typing import Any, Callable, Dict, Optional

from p2pfl_b200.management.logger import logger
from p2pfl_b200.utils.lockcheck import make_condition, make_lock, make_rlock


@dataclass
class NeighborEntry:
    link: Any  # transport-specific handle; None for non-direct
    direct: bool
    last_beat: float


class Neighbors:
    """Thread-safe address -> :class:`NeighborEntry` map."""

    def __init__(
        self,
        self_addr: str,
        open_link: Callable[[str, bool], Any],
        close_link: Callable[[str, Any, bool], None],
        on_change: Optional[Callable[[], None]] = None,
    ) -> None:
        self.self_addr = self_addr
        self._open = open_link
        self._close = close_link
        self._on_change = on_change
        self._neis: Dict[str, NeighborEntry] = {}
        self._lock = make_rlock("neighbors")

    def _changed(self) -> None:
        if self._on_change is not None:
            self._on_change()

    # ------------------------------------------------------------------ add
    def add(self, addr: str, non_direct: bool = False, handshake: bool = True) -> bool:
        """Add a neighbour; for a direct one the transport link is opened first
        (and the remote side is told via handshake unless this add *is* the
        reaction to a handshake)."""
        if addr == self.self_addr:
            logger.debug(self.self_addr, "refusing to add myself as neighbour")
            return False
        with self._lock:
            cur = self._neis.get(addr)
            if cur is not None and (cur.direct or non_direct):
                return False
        link = None
        if not non_direct:
            try:
                link = self._open(addr, handshake)
            except Exception as e:
                logger.debug(self.self_addr, f"cannot connect to {addr}: {e}")
                return False
        with self._lock:
            cur = self._neis.get(addr)
            if cur is not None and cur.direct:
                # lost a race with a concurrent add; keep the first link
                if link is not None:
                    try:
                        self._close(addr, link, False)
                    except Exception:
                        pass
                return False
            self._neis[addr] = NeighborEntry(link, not non_direct, time.time())
        self._changed()
        return True

    def refresh_or_add(self, addr: str, beat_time: float) -> None:
        """Heartbeat hook: refresh a known neighbour, learn an unknown one as
        non-direct (reference: grpc_neighbors.py:34-55)."""
        if addr == self.self_addr:
            return
        with self._lock:
            cur = self._neis.get(addr)
            if cur is not None:
                cur.last_beat = max(cur.last_beat, beat_time)
                return
        self.add(addr, non_direct=True)

    # --------------------------------------------------------------- remove
    def remove(self, addr: str, disconnect_msg: bool = True) -> None:
        with self._lock:
            entry = self._neis.pop(addr, None)
        if entry is None:
            return
        if entry.direct:
            try:
                self._close(addr, entry.link, disconnect_msg)
            except Exception:
                pass
        self._changed()

    def clear_neighbors(self) -> None:
        for addr in list(self.get_all()):
            self.remove(addr, disconnect_msg=True)

    # ---------------------------------------------------------------- query
    def get(self, addr: str) -> NeighborEntry:
        with self._lock:
            return self._neis[addr]

    def get_all(self, only_direct: bool = False) -> Dict[str, NeighborEntry]:
        with self._lock:
            if only_direct:
                return {a: e for a, e in self._neis.items() if e.direct}
            return dict(self._neis)

    def exists(self, addr: str) -> bool:
        with self._lock:
            return addr in self._neis

    def stale(self, older_than: float) -> list:
        with self._lock:
            return [a for a, e in self._neis.items() if e.last_beat < older_than]
