"""Host timeline tracing: who did what when, per node and thread, as a Chrome trace-event file.

The reference has no tracing (log lines only).  The device side of this framework is
covered by CUDA events, NVTX ranges (``Settings.NVTX_RANGES``) and ncu; what those do not show is the
HOST side of a federated round - stage changes, message handlers, relay batches, the waits between
them - which is where the time of every configuration but the single-GPU one goes.

``P2PFL_TRACE=/path/trace_{pid}.json`` (or ``tracer.enable(path)``) records spans from the stage
runner, the protocol engine (one span per handled message / model, per relay batch) and the train
stage's phases; the file is written at interpreter exit or by ``tracer.dump()`` and opens in
``chrome://tracing`` / Perfetto: one "process" per node, one track per thread.  Disabled, a span costs
one global read.
"""

from __future__ import annotations

import atexit
import json
import os
import threading
import time
from typing import Any, Dict, List, Optional, Tuple

_path: Optional[str] = os.environ.get("P2PFL_TRACE") or None
_events: List[Tuple[str, str, str, int, float, float, Optional[Dict[str, Any]]]] = []  # name, cat, node, tid, t0, dur, args
_limit = int(os.environ.get("P2PFL_TRACE_LIMIT", "2000000"))


def enabled() -> bool:
    return _path is not None


def enable(path: Optional[str]) -> None:
    """Start (``path``) or stop (``None``) recording; recorded events are kept until ``dump`` / ``reset``."""
    global _path
    _path = path


def reset() -> None:
    del _events[:]


class _Span:
    __slots__ = ("name", "cat", "node", "args", "t0")

    def __init__(self, name: str, cat: str, node: str, args: Optional[Dict[str, Any]]) -> None:
        self.name, self.cat, self.node, self.args = name, cat, node, args
        self.t0 = 0.0

    def __enter__(self) -> "_Span":
        self.t0 = time.time()
        return self

    def __exit__(self, *exc: Any) -> None:
        complete(self.name, self.node, self.t0, time.time(), self.cat, self.args)


class _Off:
    __slots__ = ()

    def __enter__(self) -> "_Off":
        return self

    def __exit__(self, *exc: Any) -> None:
        return None


_OFF = _Off()


def span(name: str, node: str, cat: str = "stage", **args: Any) -> Any:
    """``with tracer.span("TrainStage", addr, round=3): ...``"""
    if _path is None:
        return _OFF
    return _Span(name, cat, node, args or None)


def complete(name: str, node: str, t_start: float, t_end: float, cat: str = "stage", args: Optional[Dict[str, Any]] = None) -> None:
    """A span whose ends were timed by the caller (``time.time()`` values)."""
    if _path is None or len(_events) >= _limit:
        return
    _events.append((name, cat, node, threading.get_ident(), t_start, max(0.0, t_end - t_start), args))


def instant(name: str, node: str, cat: str = "mark", **args: Any) -> None:
    if _path is None or len(_events) >= _limit:
        return
    _events.append((name, cat, node, threading.get_ident(), time.time(), -1.0, args or None))


def dump(path: Optional[str] = None) -> Optional[str]:
    """Write what was recorded so far; returns the file name (None when there is nothing to write)."""
    target = path or _path
    if target is None or not _events:
        return None
    target = target.replace("{pid}", str(os.getpid()))
    events = list(_events)
    t_base = min(e[4] for e in events)
    nodes: Dict[str, int] = {}
    threads: Dict[Tuple[int, int], str] = {}
    names = {t.ident: t.name for t in threading.enumerate()}
    out: List[Dict[str, Any]] = []
    for name, cat, node, tid, t0, dur, args in events:
        pid = nodes.setdefault(node, len(nodes) + 1)
        threads.setdefault((pid, tid), names.get(tid, f"thread-{tid}"))
        ev: Dict[str, Any] = {"name": name, "cat": cat, "pid": pid, "tid": tid % 1000000, "ts": round((t0 - t_base) * 1e6, 1)}
        if dur < 0:
            ev.update(ph="i", s="t")
        else:
            ev.update(ph="X", dur=round(dur * 1e6, 1))
        if args:
            ev["args"] = {k: (v if isinstance(v, (int, float, str, bool, type(None))) else str(v)) for k, v in args.items()}
        out.append(ev)
    for node, pid in nodes.items():
        out.append({"name": "process_name", "ph": "M", "pid": pid, "args": {"name": node}})
    for (pid, tid), tname in threads.items():
        out.append({"name": "thread_name", "ph": "M", "pid": pid, "tid": tid % 1000000, "args": {"name": tname}})
    os.makedirs(os.path.dirname(os.path.abspath(target)), exist_ok=True)
    with open(target, "w") as fh:
        json.dump({"traceEvents": out, "displayTimeUnit": "ms", "otherData": {"os_pid": os.getpid(), "t_base_unix": t_base}}, fh)
    return target


def summary() -> Dict[str, Dict[str, float]]:
    """Total seconds and count per (category, name): a quick look without opening the file."""
    agg: Dict[str, Dict[str, float]] = {}
    for name, cat, _node, _tid, _t0, dur, _args in list(_events):
        if dur < 0:
            continue
        row = agg.setdefault(f"{cat}:{name}", {"seconds": 0.0, "count": 0})
        row["seconds"] += dur
        row["count"] += 1
    return agg


@atexit.register
def _dump_at_exit() -> None:
    try:
        if _path is not None:
            dump()
    except Exception:
        pass
