"""Process-wide logger / metric façade.

``logger`` is a module-level singleton with the reference's call signatures
(p2pfl/management/logger.py:144-584): ``logger.info(node, msg)``,
``log_metric(node, metric, value, step=None, round=None)``,
``get_local_logs()``, ``get_global_logs()``, ``register_node``,
``unregister_node``, ``connect_web``, ``set_level/get_level/get_level_name``.

Differences by design:
* records go through a ``QueueHandler`` so the calling (control-plane) thread
  never blocks on stdout or disk;
* ``log_metric`` for a node that lives in another process no longer raises
  ``KeyError``: the experiment/round are taken from any local node
  that is in the same experiment;
* the file handler is opt-in (``Settings.LOG_TO_FILE``).
"""

from __future__ import annotations

import atexit
import datetime
import logging
import logging.handlers
import os
import queue
import threading
from typing import Any, Dict, Optional, Tuple

from p2pfl_b200.management.metric_storage import (
    GlobalLogsType,
    GlobalMetricStorage,
    LocalLogsType,
    LocalMetricStorage,
)
from p2pfl_b200.management.node_monitor import NodeMonitor
from p2pfl_b200.settings import Settings
from p2pfl_b200.utils.lockcheck import make_condition, make_lock, make_rlock

_GREY, _YELLOW, _RED, _BOLD_RED, _CYAN, _RESET = (
    "\033[90m",
    "\033[93m",
    "\033[91m",
    "\033[1;91m",
    "\033[96m",
    "\033[0m",
)
_LEVEL_COLOR = {
    logging.DEBUG: _GREY,
    logging.INFO: _CYAN,
    logging.WARNING: _YELLOW,
    logging.ERROR: _RED,
    logging.CRITICAL: _BOLD_RED,
}


class _ConsoleFormatter(logging.Formatter):
    def format(self, record: logging.LogRecord) -> str:
        ts = datetime.datetime.fromtimestamp(record.created).strftime("%H:%M:%S.%f")[:-3]
        color = _LEVEL_COLOR.get(record.levelno, "")
        node = getattr(record, "node", "-")
        return f"{_GREY}{ts}{_RESET} {color}{record.levelname:<8}{_RESET} [{node}] {record.getMessage()}"


class _PlainFormatter(logging.Formatter):
    def format(self, record: logging.LogRecord) -> str:
        ts = datetime.datetime.fromtimestamp(record.created).isoformat(timespec="milliseconds")
        return f"{ts} {record.levelname} [{getattr(record, 'node', '-')}] {record.getMessage()}"


class _WebHandler(logging.Handler):
    """Forwards records to the dashboard (reference: logger.py:68-113)."""

    def __init__(self, web: Any) -> None:
        super().__init__()
        self._web = web

    def emit(self, record: logging.LogRecord) -> None:
        try:
            self._web.send_log(
                datetime.datetime.fromtimestamp(record.created),
                getattr(record, "node", "-"),
                record.levelno,
                record.getMessage(),
            )
        except Exception:
            pass  # never let observability break the node


class Logger:
    """Singleton façade; use the module-level ``logger`` instance."""

    def __init__(self) -> None:
        self._log = logging.getLogger("p2pfl_b200")
        self._log.propagate = False
        self._log.setLevel(getattr(logging, Settings.LOG_LEVEL, logging.INFO))
        self._queue: "queue.SimpleQueue[Any]" = queue.SimpleQueue()
        self._log.handlers = [logging.handlers.QueueHandler(self._queue)]
        console = logging.StreamHandler()
        console.setFormatter(_ConsoleFormatter())
        self._sinks = [console]
        if Settings.LOG_TO_FILE:
            self._add_file_sink()
        self._listener = logging.handlers.QueueListener(self._queue, *self._sinks)
        self._listener.start()
        atexit.register(self.cleanup)

        self._web: Any = None
        self._nodes: Dict[str, Tuple[Optional[NodeMonitor], Any]] = {}
        self._nodes_lock = make_lock("logger.nodes")
        self.local_metrics = LocalMetricStorage()
        self.global_metrics = GlobalMetricStorage()

    # ------------------------------------------------------------- plumbing
    def _add_file_sink(self) -> None:
        os.makedirs(Settings.LOG_DIR, exist_ok=True)
        fh = logging.handlers.RotatingFileHandler(
            os.path.join(Settings.LOG_DIR, "p2pfl.log"), maxBytes=1_000_000, backupCount=3
        )
        fh.setFormatter(_PlainFormatter())
        self._sinks.append(fh)

    def _restart_listener(self) -> None:
        self._listener.stop()
        self._listener = logging.handlers.QueueListener(self._queue, *self._sinks)
        self._listener.start()

    def cleanup(self) -> None:
        """Stop monitors and flush the queue (reference: logger.py:234-249)."""
        with self._nodes_lock:
            for mon, _ in self._nodes.values():
                if mon is not None:
                    mon.stop()
            self._nodes.clear()
        try:
            self._listener.stop()
        except Exception:
            pass

    def connect_web(self, url: str, key: str, session: Any = None) -> None:
        """Attach the REST dashboard (reference: logger.py:163-179)."""
        from p2pfl_b200.management.web_services import P2pflWebServices

        self._web = P2pflWebServices(url, key, session=session)
        self._sinks.append(_WebHandler(self._web))
        self._restart_listener()

    # ---------------------------------------------------------------- level
    def set_level(self, level: Any) -> None:
        if isinstance(level, str):
            level = getattr(logging, level.upper())
        self._log.setLevel(level)

    def get_level(self) -> int:
        return self._log.getEffectiveLevel()

    @staticmethod
    def get_level_name(lvl: int) -> str:
        return logging.getLevelName(lvl)

    # -------------------------------------------------------------- logging
    def log(self, level: int, node: str, message: str) -> None:
        if self._log.isEnabledFor(level):
            self._log.log(level, message, extra={"node": node})

    def info(self, node: str, message: str) -> None:
        self.log(logging.INFO, node, message)

    def debug(self, node: str, message: str) -> None:
        self.log(logging.DEBUG, node, message)

    def warning(self, node: str, message: str) -> None:
        self.log(logging.WARNING, node, message)

    def error(self, node: str, message: str) -> None:
        self.log(logging.ERROR, node, message)

    def critical(self, node: str, message: str) -> None:
        self.log(logging.CRITICAL, node, message)

    def is_debug(self) -> bool:
        return self._log.isEnabledFor(logging.DEBUG)

    # -------------------------------------------------------------- metrics
    def _experiment_of(self, node: str) -> Tuple[Optional[str], Optional[int]]:
        with self._nodes_lock:
            entry = self._nodes.get(node)
            if entry is not None and entry[1].actual_exp_name is not None:
                return entry[1].actual_exp_name, entry[1].round
            # remote source: borrow the coordinates of a local node that is learning
            for _, st in self._nodes.values():
                if st.actual_exp_name is not None:
                    return st.actual_exp_name, st.round
        return None, None

    def log_metric(
        self, node: str, metric: str, value: float, step: Optional[int] = None, round: Optional[int] = None
    ) -> None:
        """``step`` given -> local (training) metric; otherwise global (per round)
        (reference: logger.py:393-438)."""
        exp, cur_round = self._experiment_of(node)
        if exp is None:
            exp = "experiment"
        if round is None:
            round = cur_round if cur_round is not None else 0
        if step is None:
            self.global_metrics.add_log(exp, round, metric, node, value)
            if self._web is not None:
                try:
                    self._web.send_global_metric(exp, round, metric, node, value)
                except Exception:
                    pass
        else:
            self.local_metrics.add_log(exp, round, metric, node, value, step)
            if self._web is not None:
                try:
                    self._web.send_local_metric(exp, round, metric, node, value, step)
                except Exception:
                    pass

    def log_system_metric(self, node: str, metric: str, value: float, time: Optional[datetime.datetime] = None) -> None:
        if self._web is not None:
            try:
                self._web.send_system_metric(node, metric, value, time or datetime.datetime.now())
            except Exception:
                pass

    def get_local_logs(self) -> LocalLogsType:
        return self.local_metrics.get_all_logs()

    def get_global_logs(self) -> GlobalLogsType:
        return self.global_metrics.get_all_logs()

    def reset_metrics(self) -> None:
        self.local_metrics = LocalMetricStorage()
        self.global_metrics = GlobalMetricStorage()

    # ------------------------------------------------------------- registry
    def register_node(self, node: str, state: Any, simulation: bool = False, gpu_index: Optional[int] = None) -> None:
        """reference: logger.py:492-519.  A monitor thread is only started when
        the dashboard is connected, as in the reference (logger.py:505-511)."""
        with self._nodes_lock:
            if node in self._nodes:
                raise ValueError(f"node {node} already registered")
            mon: Optional[NodeMonitor] = None
            if self._web is not None:
                try:
                    self._web.register_node(node, simulation)
                except Exception:
                    pass
                mon = NodeMonitor(node, lambda n, k, v: self.log_system_metric(n, k, v), gpu_index)
                mon.start()
            self._nodes[node] = (mon, state)

    def unregister_node(self, node: str) -> None:
        with self._nodes_lock:
            entry = self._nodes.pop(node, None)
        if entry is None:
            return
        if entry[0] is not None:
            entry[0].stop()
        if self._web is not None:
            self._web.unregister_node(node)

    def registered_nodes(self) -> Dict[str, Any]:
        with self._nodes_lock:
            return {k: v[1] for k, v in self._nodes.items()}

    # ---------------------------------------------- lifecycle notifications
    def experiment_started(self, node: str) -> None:
        self.debug(node, "experiment started")

    def experiment_finished(self, node: str) -> None:
        self.debug(node, "experiment finished")

    def round_finished(self, node: str) -> None:
        self.debug(node, "round finished")


logger = Logger()
