"""Heartbeats and failure detection.

Every ``HEARTBEAT_PERIOD`` the thread evicts neighbours whose last beat is
older than ``HEARTBEAT_TIMEOUT`` and broadcasts ``beat [now]`` to the direct
neighbours; beats are TTL-flooded by the receivers, which is also how
non-direct nodes are discovered (reference: p2pfl/communication/heartbeater.py:33-111).
"""

from __future__ import annotations

import threading
import time
from typing import Callable

from p2pfl_b200.communication.neighbors import Neighbors
from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings

HEARTBEAT_CMD = "beat"


class Heartbeater(threading.Thread):
    def __init__(self, self_addr: str, neighbors: Neighbors, broadcast_beat: Callable[[float], None]) -> None:
        super().__init__(name=f"heartbeater-thread-{self_addr}", daemon=True)
        self._addr = self_addr
        self._neighbors = neighbors
        self._broadcast_beat = broadcast_beat
        self._stop_evt = threading.Event()

    def stop(self) -> None:
        self._stop_evt.set()

    def beat(self, nei: str, beat_time: float) -> None:
        """Inbound beat from ``nei`` (direct or relayed)."""
        if nei == self._addr:
            return
        self._neighbors.refresh_or_add(nei, beat_time)

    def run(self) -> None:
        while not self._stop_evt.is_set():
            t0 = time.time()
            for nei in self._neighbors.stale(t0 - Settings.HEARTBEAT_TIMEOUT):
                logger.info(self._addr, f"heartbeat timeout for {nei}, removing it")
                self._neighbors.remove(nei, disconnect_msg=False)
            try:
                self._broadcast_beat(t0)
            except Exception as e:
                logger.debug(self._addr, f"beat broadcast failed: {e}")
            self._stop_evt.wait(max(0.0, Settings.HEARTBEAT_PERIOD - (time.time() - t0)))
