"""In-memory metric stores.

Two shapes, as in the reference (p2pfl/management/metric_storage.py:30-247):

* local  (per-step training metrics):  exp -> round -> node -> metric -> [(step, value)]
* global (per-round evaluation metrics): exp -> node -> metric -> [(round, value)]
  with at most one value per (metric, round).

Both are lock-protected and return deep-ish copies so callers can iterate
without holding the lock.
"""

from __future__ import annotations

import threading
from typing import Dict, List, Tuple, Union
from p2pfl_b200.utils.lockcheck import make_condition, make_lock, make_rlock

Number = Union[int, float]
MetricsType = Dict[str, List[Tuple[int, float]]]
NodeLogsType = Dict[str, MetricsType]
RoundLogsType = Dict[int, NodeLogsType]
LocalLogsType = Dict[str, RoundLogsType]
GlobalLogsType = Dict[str, NodeLogsType]


def _copy_metrics(m: MetricsType) -> MetricsType:
    return {k: list(v) for k, v in m.items()}


class LocalMetricStorage:
    """Per-step metrics produced while a node trains."""

    def __init__(self) -> None:
        self._logs: LocalLogsType = {}
        self._lock = make_lock("metrics")

    def add_log(self, exp_name: str, round: int, metric: str, node: str, val: Number, step: int) -> None:
        with self._lock:
            series = (
                self._logs.setdefault(exp_name, {})
                .setdefault(round, {})
                .setdefault(node, {})
                .setdefault(metric, [])
            )
            series.append((step, float(val)))

    def get_all_logs(self) -> LocalLogsType:
        with self._lock:
            return {
                e: {r: {n: _copy_metrics(m) for n, m in nodes.items()} for r, nodes in rounds.items()}
                for e, rounds in self._logs.items()
            }

    def get_experiment_logs(self, exp: str) -> RoundLogsType:
        return self.get_all_logs()[exp]

    def get_experiment_round_logs(self, exp: str, round: int) -> NodeLogsType:
        return self.get_experiment_logs(exp)[round]

    def get_experiment_round_node_logs(self, exp: str, round: int, node: str) -> MetricsType:
        return self.get_experiment_round_logs(exp, round)[node]


class GlobalMetricStorage:
    """Per-round metrics (evaluation results, shared between nodes)."""

    def __init__(self) -> None:
        self._logs: GlobalLogsType = {}
        self._lock = make_lock("metrics")

    def add_log(self, exp_name: str, round: int, metric: str, node: str, val: Number) -> None:
        with self._lock:
            series = self._logs.setdefault(exp_name, {}).setdefault(node, {}).setdefault(metric, [])
            if all(r != round for r, _ in series):
                series.append((round, float(val)))

    def get_all_logs(self) -> GlobalLogsType:
        with self._lock:
            return {e: {n: _copy_metrics(m) for n, m in nodes.items()} for e, nodes in self._logs.items()}

    def get_experiment_logs(self, exp: str) -> NodeLogsType:
        return self.get_all_logs()[exp]

    def get_experiment_node_logs(self, exp: str, node: str) -> MetricsType:
        return self.get_experiment_logs(exp)[node]
