"""Test / simulation helpers (reference: p2pfl/utils.py:37-138)."""

from __future__ import annotations

import time
from typing import Any, List

import numpy as np

from p2pfl_b200.settings import Settings, set_test_settings

__all__ = ["set_test_settings", "wait_convergence", "full_connection", "wait_4_results", "check_equal_models"]


def wait_convergence(nodes: List[Any], n_neis: int, wait: float = 5, only_direct: bool = False) -> None:
    """Block until every node sees exactly ``n_neis`` neighbours; raise after ``wait`` seconds."""
    deadline = time.time() + wait
    while True:
        if all(len(n.get_neighbors(only_direct=only_direct)) == n_neis for n in nodes):
            return
        if time.time() > deadline:
            seen = [len(n.get_neighbors(only_direct=only_direct)) for n in nodes]
            raise AssertionError(f"neighbour convergence not reached: want {n_neis}, have {seen}")
        time.sleep(0.02)


def full_connection(node: Any, nodes: List[Any]) -> None:
    """Connect ``node`` to each of ``nodes``."""
    for n in nodes:
        node.connect(n.addr)


def wait_4_results(nodes: List[Any], timeout: float = 600.0) -> None:
    """Block until every node finished its experiment (``state.round is None``)."""
    deadline = time.time() + timeout
    for n in nodes:
        left = deadline - time.time()
        if left <= 0 or not n.wait_learning_finished(left):
            raise TimeoutError(f"node {n.addr} did not finish learning in {timeout}s")


def check_equal_models(nodes: List[Any], atol: float = 1e-1) -> None:
    """All nodes hold the same parameters (reference tolerance is 1e-1, Q15;
    callers here pass far tighter ones)."""
    ref = None
    for node in nodes:
        learner = node.state.learner
        if learner is None:
            raise AssertionError(f"node {node.addr} has no learner")
        params = {k: np.asarray(v.detach().float().cpu()) for k, v in learner.get_parameters().items()}
        if ref is None:
            ref = params
            continue
        for k in ref:
            if not np.allclose(ref[k], params[k], atol=atol):
                diff = float(np.max(np.abs(ref[k] - params[k])))
                raise AssertionError(f"models differ at '{k}' (max abs diff {diff})")
