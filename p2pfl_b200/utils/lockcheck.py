"""Lock-order checking for the control plane (a poor man's ThreadSanitizer for lock cycles).

The reference has no race tooling at all and its threads - learning thread, gossiper,
heartbeater, one RPC handler per inbound message - share per-node state under a handful of locks.
A deadlock needs two locks taken in opposite orders by two threads; that can be detected without
ever deadlocking: record, for every acquisition, which locks the acquiring thread already holds, and
keep the union of those "held -> acquired" edges.  A cycle in that graph is an ordering two threads
can realise in opposite directions.

``make_lock(name)`` / ``make_rlock(name)`` / ``make_condition(name)`` return plain ``threading``
objects unless checking is on (``P2PFL_LOCKCHECK=1`` or ``lockcheck.enable()``), in which case they
are wrapped.  Locks are identified by NAME (all ``neighbors`` locks of all nodes are one vertex): the
rule being checked is the code's ordering discipline, not an instance-level history.  Nesting two
locks of the SAME name is not an edge (re-entrant locks, and the in-memory transport, where node A's
handler calls straight into node B's handler and so nests the same kinds of locks of two nodes).

``violations()`` returns the cycles found so far; the stress tests assert it stays empty.
"""

from __future__ import annotations

import os
import threading
from typing import Any, Dict, List, Optional, Set, Tuple

_enabled = os.environ.get("P2PFL_LOCKCHECK", "0") not in ("", "0")
_graph: Dict[str, Set[str]] = {}
_graph_lock = threading.Lock()
_violations: List[Tuple[str, ...]] = []
_tls = threading.local()


def enable(on: bool = True) -> None:
    """Wrap locks created FROM NOW ON (existing objects keep what they have)."""
    global _enabled
    _enabled = on


def enabled() -> bool:
    return _enabled


def reset() -> None:
    with _graph_lock:
        _graph.clear()
        _violations.clear()


def violations() -> List[Tuple[str, ...]]:
    with _graph_lock:
        return list(_violations)


def edges() -> Dict[str, Set[str]]:
    with _graph_lock:
        return {k: set(v) for k, v in _graph.items()}


def _held() -> List[str]:
    h = getattr(_tls, "held", None)
    if h is None:
        h = _tls.held = []
    return h


def _path(src: str, dst: str) -> Optional[List[str]]:
    """A path src -> ... -> dst in the order graph (caller holds _graph_lock)."""
    stack, seen = [(src, [src])], {src}
    while stack:
        node, path = stack.pop()
        if node == dst:
            return path
        for nxt in _graph.get(node, ()):
            if nxt not in seen:
                seen.add(nxt)
                stack.append((nxt, path + [nxt]))
    return None


def _note_acquire(name: str) -> None:
    held = _held()
    if held:
        with _graph_lock:
            for h in dict.fromkeys(held):  # distinct, in order
                if h == name:
                    continue  # re-entrant / same-kind nesting: see module docstring
                if name not in _graph.get(h, ()):
                    back = _path(name, h)  # would h -> name close a cycle?
                    if back is not None:
                        cycle = tuple(back + [name])
                        if cycle not in _violations:
                            _violations.append(cycle)
                    _graph.setdefault(h, set()).add(name)
    held.append(name)


def _note_release(name: str) -> None:
    held = _held()
    for i in range(len(held) - 1, -1, -1):
        if held[i] == name:
            del held[i]
            return


class _Tracked:
    """Context-manager / acquire-release facade over a threading lock that reports to the order graph."""

    def __init__(self, inner: Any, name: str) -> None:
        self._inner, self._name = inner, name

    def acquire(self, blocking: bool = True, timeout: float = -1) -> bool:
        ok = self._inner.acquire(blocking, timeout)
        if ok:
            _note_acquire(self._name)
        return ok

    def release(self) -> None:
        _note_release(self._name)
        self._inner.release()

    def locked(self) -> bool:
        return self._inner.locked()

    def __enter__(self) -> "_Tracked":
        self.acquire()
        return self

    def __exit__(self, *exc: Any) -> None:
        self.release()


class _TrackedCondition(threading.Condition):
    """``threading.Condition`` whose ``with`` blocks are reported (``threading.Condition`` binds its
    ``acquire`` / ``release`` per instance, so only the context-manager form is tracked; a thread parked in
    ``wait`` keeps the name on its list, but it acquires nothing while it sleeps)."""

    def __init__(self, name: str) -> None:
        super().__init__(threading.RLock())
        self._lc_name = name

    def __enter__(self) -> bool:
        r = super().__enter__()
        _note_acquire(self._lc_name)
        return r

    def __exit__(self, *exc: Any) -> None:
        _note_release(self._lc_name)
        return super().__exit__(*exc)


def make_lock(name: str) -> Any:
    return _Tracked(threading.Lock(), name) if _enabled else threading.Lock()


def make_rlock(name: str) -> Any:
    return _Tracked(threading.RLock(), name) if _enabled else threading.RLock()


def make_condition(name: str) -> Any:
    return _TrackedCondition(name) if _enabled else threading.Condition()
