"""Aggregation bookkeeping (host side).

Admission rules are the reference's (p2pfl/learning/aggregators/aggregator.py:117-200):

* *waiting mode* (non-trainers): only a model whose contributors equal the
  whole train set is accepted;
* *normal mode* (trainers): a full-train-set aggregate replaces everything;
  a partial is accepted only if all its contributors are in the train set and
  disjoint from what is already held; anything else is rejected (returns ``[]``).

A held entry is keyed by the space-joined contributor list and stores the
*normalised* partial mean with its total weight, so re-aggregation is exact
(SURVEY Appendix B).  The reference's lock-as-latch is an ``Event`` here and
an empty contributor list is rejected without touching any lock (Q3).

The arithmetic itself is delegated to :meth:`aggregate`; models are opaque
(state-dicts on the host path, ``FlatParams`` on the device path).
"""

from __future__ import annotations

import threading
from typing import Any, Dict, List, Optional, Tuple

from p2pfl_b200.exceptions import NoModelsToAggregateError
from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings
from p2pfl_b200.utils.lockcheck import make_condition, make_lock, make_rlock

ModelEntry = Tuple[Any, int]


class Aggregator:
    """Base class; subclasses implement :meth:`aggregate`."""

    def __init__(self, node_name: str = "unknown") -> None:
        self.node_name = node_name
        self._train_set: List[str] = []
        self._waiting = False
        self._models: Dict[str, ModelEntry] = {}
        self._lock = make_rlock("aggregator")
        self._done = threading.Event()
        self._done.set()  # nothing pending
        self._listeners: List[Any] = []

    # ------------------------------------------------------------- strategy
    def aggregate(self, models: Dict[str, ModelEntry]) -> Any:
        raise NotImplementedError

    # ---------------------------------------------------------------- setup
    def set_nodes_to_aggregate(self, nodes: List[str]) -> None:
        with self._lock:
            if self._train_set and not self._done.is_set():
                raise RuntimeError("an aggregation is already in progress; call clear() first")
            self._train_set = list(nodes)
            self._waiting = False
            self._models = {}
            self._done.clear()

    def set_waiting_aggregated_model(self, nodes: List[str]) -> None:
        self.set_nodes_to_aggregate(nodes)
        with self._lock:
            self._waiting = True

    def clear(self) -> None:
        with self._lock:
            self._train_set = []
            self._waiting = False
            self._models = {}
            self._done.set()

    def add_listener(self, fn: Any) -> None:
        """``fn()`` is called after every accepted model (used to wake gossip loops)."""
        self._listeners.append(fn)

    # ---------------------------------------------------------------- query
    def get_train_set(self) -> List[str]:
        with self._lock:
            return list(self._train_set)

    def is_waiting(self) -> bool:
        return self._waiting

    def get_aggregated_models(self) -> List[str]:
        """Flat list of every contributor currently held."""
        with self._lock:
            out: List[str] = []
            for key in self._models:
                out.extend(key.split())
            return out

    # ------------------------------------------------------------------ add
    def add_model(self, model: Any, contributors: List[str], weight: int) -> List[str]:
        """Offer a model; returns the contributors now held, or ``[]`` if it
        was rejected."""
        if not contributors:
            logger.debug(self.node_name, "model without contributors rejected")
            return []
        accepted = False
        with self._lock:
            if not self._train_set or self._done.is_set():
                logger.debug(self.node_name, "received a model when none is needed")
                return []
            full = set(contributors) == set(self._train_set)
            if self._waiting:
                if not full:
                    logger.debug(self.node_name, "waiting for the full aggregate; partial ignored")
                    return []
                self._models = {" ".join(contributors): (model, 1)}
                accepted = True
                self._done.set()
            else:
                held = set(self.get_aggregated_models())
                if not all(c in self._train_set for c in contributors):
                    logger.debug(self.node_name, f"model from outside the train set rejected: {contributors}")
                    return []
                if full:
                    self._models = {" ".join(contributors): (model, weight)}
                    accepted = True
                    self._done.set()
                elif held.isdisjoint(contributors):
                    self._models[" ".join(contributors)] = (model, weight)
                    accepted = True
                    if len(held) + len(contributors) >= len(self._train_set):
                        self._done.set()
                else:
                    logger.debug(self.node_name, f"overlapping partial rejected: {contributors} vs {sorted(held)}")
                    return []
            now_held = self.get_aggregated_models()
        if accepted:
            if logger.is_debug():
                logger.debug(self.node_name, f"model added ({len(now_held)}/{len(self._train_set)}) from {contributors}")
            for fn in self._listeners:
                try:
                    fn()
                except Exception:
                    pass
        return now_held

    def would_accept(self, contributors: List[str]) -> bool:
        """``add_model``'s admission rules without the model: lets a receiver skip decoding (a pull over
        NVLink on the device path) a copy it has no use for - in a mesh every holder of the aggregate
        offers it to every neighbour that has not announced ``models_ready`` yet, so a waiting node is
        offered the same model by most of its neighbours within a millisecond."""
        if not contributors:
            return False
        with self._lock:
            if not self._train_set or self._done.is_set():
                return False
            full = set(contributors) == set(self._train_set)
            if self._waiting:
                return full
            if not all(c in self._train_set for c in contributors):
                return False
            return full or set(self.get_aggregated_models()).isdisjoint(contributors)

    def remove_node(self, node: str) -> None:
        """A train-set member vanished: stop expecting its model."""
        with self._lock:
            if node in self._train_set and node not in self.get_aggregated_models() and not self._done.is_set():
                self._train_set = [n for n in self._train_set if n != node]
                if not self._waiting and len(self.get_aggregated_models()) >= len(self._train_set) > 0:
                    self._done.set()

    # ------------------------------------------------------------------ get
    def wait_and_get_aggregation(self, timeout: Optional[float] = None) -> Any:
        """Block until complete (or timeout) and return the aggregate of what
        is held (reference: aggregator.py:202-247)."""
        if timeout is None:
            timeout = Settings.AGGREGATION_TIMEOUT
        finished = self._done.wait(timeout)
        with self._lock:
            models = dict(self._models)
            train_set = list(self._train_set)
            waiting = self._waiting
        if not finished:
            missing = set(train_set) - set(c for k in models for c in k.split())
            logger.info(self.node_name, f"aggregation timeout; missing models: {sorted(missing)}")
        if not models:
            raise NoModelsToAggregateError(f"({self.node_name}) no models to aggregate")
        if waiting or len(models) == 1:
            # a single entry is already the normalised mean of its contributors
            return next(iter(models.values()))[0]
        return self.aggregate(models)

    def get_partial_aggregation(self, except_nodes: List[str]) -> Tuple[Any, Optional[List[str]], Optional[int]]:
        """Aggregate every held entry that shares no contributor with
        ``except_nodes`` (reference: aggregator.py:249-281)."""
        with self._lock:
            chosen = {
                k: v for k, v in self._models.items() if not any(c in except_nodes for c in k.split())
            }
        if not chosen:
            return None, None, None
        contributors: List[str] = []
        for k in chosen:
            contributors.extend(k.split())
        total = sum(w for _, w in chosen.values())
        if len(chosen) == 1:
            return next(iter(chosen.values()))[0], contributors, total
        return self.aggregate(chosen), contributors, total
