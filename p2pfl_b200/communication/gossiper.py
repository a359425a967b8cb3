"""Gossip engine.

Two jobs, as in the reference's ``Gossiper`` (p2pfl/communication/gossiper.py:31-243):

(a) relay of small control messages: a rate-limited queue drained by one
    daemon thread, plus the duplicate filter (ring of recent message ids);
(b) ``gossip_weights``: the sender-driven model push loop that runs on the
    caller's (learning) thread.

Design changes: the relay thread sleeps on a condition and wakes the moment a
message is queued instead of polling every ``GOSSIP_PERIOD``; the model loop
waits on an event the caller supplies so a state change (an ack from a peer)
re-evaluates candidates immediately, with ``period`` only as the upper bound;
pacing uses ``period - elapsed`` (Q4); stagnation needs *all* of the last X
statuses equal (Q5); the thread is a daemon (Q12).
"""

from __future__ import annotations

import collections
import random
import threading
import time
from typing import Any, Callable, Deque, List, Optional, Set, Tuple

from p2pfl_b200.communication.messages import Message
from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings
from p2pfl_b200.utils.lockcheck import make_condition, make_lock, make_rlock

SendFn = Callable[[str, Any, bool], None]


class Gossiper(threading.Thread):
    def __init__(self, self_addr: str, send: SendFn) -> None:
        super().__init__(name=f"gossiper-thread-{self_addr}", daemon=True)
        self._addr = self_addr
        self._send = send
        self._pending: Deque[Tuple[Message, List[str]]] = collections.deque()
        self._cv = make_condition("gossiper.queue")
        self._stop_flag = False
        # lazily queued broadcasts wait here (at most BUNDLE_LINGER) for the next ordinary one
        self._lazy: List[Tuple[Message, Any]] = []
        self._lazy_deadline = 0.0
        self.bundle_fn: Optional[Callable[[List[Message]], None]] = None
        # several relayed messages for the same neighbour in one drain -> one message (set by the engine)
        self.bundle_to_fn: Optional[Callable[[str, List[Message]], None]] = None
        # duplicate filter: ring + set for O(1) membership
        self._seen_ring: Deque[int] = collections.deque()
        self._seen: Set[int] = set()
        self._seen_lock = make_lock("gossiper.seen")

    # ------------------------------------------------------------ lifecycle
    def stop(self) -> None:
        with self._cv:
            self._stop_flag = True
            self._cv.notify_all()

    # ------------------------------------------------------ duplicate filter
    def check_and_set_processed(self, msg_hash: int) -> bool:
        """True if the id is new (and is now recorded), False for a duplicate."""
        with self._seen_lock:
            if msg_hash in self._seen:
                return False
            self._seen.add(msg_hash)
            self._seen_ring.append(msg_hash)
            while len(self._seen_ring) > max(1, Settings.AMOUNT_LAST_MESSAGES_SAVED):
                self._seen.discard(self._seen_ring.popleft())
            return True

    # ---------------------------------------------------------- relay queue
    def add_broadcast(self, msg: Message, broadcast_fn: Callable[[Message], None], lazy: bool = False) -> None:
        """Queue a whole broadcast (performed with overlapped round trips by ``broadcast_fn``).
        ``lazy``: nobody is waiting for it - hold it until the next ordinary broadcast (or
        ``BUNDLE_LINGER``) so that both can travel in one message."""
        with self._cv:
            if lazy and self.bundle_fn is not None and Settings.BUNDLE_BOOKKEEPING:
                if not self._lazy:
                    self._lazy_deadline = time.time() + Settings.BUNDLE_LINGER
                self._lazy.append((msg, broadcast_fn))
                self._cv.notify()  # the relay thread re-arms its timeout
                return
            self._flush_lazy_locked()
            self._pending.append((msg, broadcast_fn))  # type: ignore[arg-type]
            self._cv.notify()

    def _flush_lazy_locked(self) -> None:
        if self._lazy:
            self._pending.extend(self._lazy)  # type: ignore[arg-type]
            self._lazy = []

    def add_message(self, msg: Message, pending_neis: List[str]) -> None:
        if not pending_neis:
            return
        with self._cv:
            self._pending.append((msg, list(pending_neis)))
            self._cv.notify()

    def run(self) -> None:
        try:  # this thread sends models, which may stage them on the device
            from p2pfl_b200 import ops

            ops.thread_uses_cuda()
        except Exception:
            pass
        while True:
            with self._cv:
                while not self._pending and not self._stop_flag:
                    if self._lazy:
                        rest = self._lazy_deadline - time.time()
                        if rest <= 0:
                            self._flush_lazy_locked()
                            break
                        self._cv.wait(rest)
                    else:
                        self._cv.wait()
                if self._stop_flag:
                    return
                budget = max(1, Settings.GOSSIP_MESSAGES_PER_PERIOD)
                batch: List[Tuple[Message, str]] = []
                while self._pending and budget > 0:
                    msg, neis = self._pending[0]
                    if callable(neis):  # a queued broadcast
                        self._pending.popleft()
                        batch.append((msg, neis))  # type: ignore[arg-type]
                        budget -= 1
                        continue
                    take = neis[:budget]
                    del neis[: len(take)]
                    budget -= len(take)
                    batch.extend((msg, n) for n in take)
                    if not neis:
                        self._pending.popleft()
            t0 = time.time()
            bcasts = [m for m, n in batch if callable(n)]
            if len(bcasts) >= 2 and self.bundle_fn is not None and Settings.BUNDLE_BOOKKEEPING:
                # several whole broadcasts drained together: one message per neighbour
                try:
                    self.bundle_fn(bcasts)
                except Exception as e:
                    logger.debug(self._addr, f"bundled broadcast failed: {e}")
                batch = [(m, n) for m, n in batch if not callable(n)]
            if self.bundle_to_fn is not None and Settings.BUNDLE_BOOKKEEPING:
                # relays: everything this drain holds for one neighbour travels together (on a line or
                # a ring a node forwards nearly every broadcast to the same one or two neighbours)
                per_nei: dict = {}
                for msg, nei in batch:
                    if not callable(nei):
                        per_nei.setdefault(nei, []).append(msg)
                grouped = {n for n, ms in per_nei.items() if len(ms) >= 2}
                for nei in grouped:
                    try:
                        self.bundle_to_fn(nei, per_nei[nei])
                    except Exception as e:
                        logger.debug(self._addr, f"bundled relay to {nei} failed: {e}")
                if grouped:
                    batch = [(m, n) for m, n in batch if callable(n) or n not in grouped]
            for msg, nei in batch:
                try:
                    if callable(nei):
                        nei(msg)
                        continue
                    self._send(nei, msg, False)
                except Exception as e:  # a dead neighbour must not kill the relay thread
                    logger.debug(self._addr, f"relay to {nei} failed: {e}")
            rest = Settings.GOSSIP_PERIOD - (time.time() - t0)
            if rest > 0 and self._pending:
                time.sleep(rest)

    # ---------------------------------------------------------- model gossip
    def gossip_weights(
        self,
        early_stopping_fn: Callable[[], bool],
        get_candidates_fn: Callable[[], List[str]],
        status_fn: Callable[[], Any],
        model_fn: Callable[[str], Any],
        period: Optional[float] = None,
        create_connection: bool = False,
        wait_fn: Optional[Callable[[float], None]] = None,
    ) -> None:
        """Push models until nobody needs one (reference: gossiper.py:167-243).

        ``wait_fn(timeout)`` blocks until something relevant changed or the
        timeout passed; default is a plain sleep."""
        if period is None:
            period = Settings.GOSSIP_MODELS_PERIOD
        window = max(1, Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS)
        last_status: Deque[str] = collections.deque(maxlen=window)
        while True:
            t0 = time.time()
            if early_stopping_fn():
                logger.debug(self._addr, "model gossip stopped early")
                return
            candidates = get_candidates_fn()
            if not candidates:
                logger.debug(self._addr, "model gossip finished: no candidates left")
                return
            last_status.append(repr(status_fn()))
            if len(last_status) == window and len(set(last_status)) == 1:
                logger.info(self._addr, f"model gossip stagnated after {window} equal ticks, leaving")
                return
            picked = random.sample(candidates, min(Settings.GOSSIP_MODELS_PER_ROUND, len(candidates)))
            for nei in picked:
                model = model_fn(nei)
                if model is None:
                    continue
                try:
                    self._send(nei, model, create_connection)
                except Exception as e:
                    logger.debug(self._addr, f"model send to {nei} failed: {e}")
            rest = period - (time.time() - t0)
            if rest > 0:
                if wait_fn is not None:
                    wait_fn(rest)
                else:
                    time.sleep(rest)
