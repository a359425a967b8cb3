"""REST client for the optional p2pfl-web dashboard.

Endpoint paths and JSON bodies follow the reference client
(p2pfl/management/p2pfl_web_services.py:82-265): ``POST /node``, ``/node-log``,
``/node-metric/local``, ``/node-metric/global``, ``/node-metric/system`` with an
``x-api-key`` header.  There is no network in the build sandbox, so this module
is exercised in tests against a stub ``session`` object only.
"""

from __future__ import annotations

import datetime
from typing import Any, Dict, Optional


class P2pflWebServicesError(Exception):
    """Non-2xx answer from the dashboard."""

    def __init__(self, code: int, message: str) -> None:
        super().__init__(f"web services error {code}: {message}")
        self.code = code
        self.message = message


class P2pflWebServices:
    """Thin, synchronous REST client.  ``session`` needs a ``post`` method with the
    ``requests`` signature; it defaults to ``requests`` itself."""

    def __init__(self, url: str, key: str, session: Any = None) -> None:
        self._url = url.rstrip("/")
        self._key = key
        self._node_ids: Dict[str, int] = {}
        if session is None:
            import requests

            session = requests
        self._session = session

    def _headers(self) -> Dict[str, str]:
        return {"Content-Type": "application/json", "x-api-key": self._key}

    def _post(self, path: str, body: Dict[str, Any]) -> Any:
        resp = self._session.post(self._url + path, json=body, headers=self._headers(), timeout=5)
        if not (200 <= resp.status_code < 300):
            raise P2pflWebServicesError(resp.status_code, getattr(resp, "text", ""))
        return resp

    def _node_id(self, node: str) -> int:
        if node not in self._node_ids:
            raise ValueError(f"node {node} is not registered with the web services")
        return self._node_ids[node]

    # -- API -----------------------------------------------------------------
    def register_node(self, node: str, is_simulated: bool) -> None:
        resp = self._post("/node", {"address": node, "is_simulated": is_simulated})
        self._node_ids[node] = resp.json()["node_id"]

    def unregister_node(self, node: str) -> None:
        # The reference leaves this unimplemented (p2pfl_web_services.py:106-114);
        # forgetting the id locally is enough for the client side.
        self._node_ids.pop(node, None)

    def send_log(self, time: datetime.datetime, node: str, level: int, message: str) -> None:
        self._post(
            "/node-log",
            {
                "time": time.strftime("%Y-%m-%d %H:%M:%S"),
                "node_id": self._node_id(node),
                "level": level,
                "message": message,
            },
        )

    def send_local_metric(self, exp: str, round: int, metric: str, node: str, value: float, step: int) -> None:
        self._post(
            "/node-metric/local",
            {
                "node_id": self._node_id(node),
                "experiment": exp,
                "round": round,
                "metric": metric,
                "value": value,
                "step": step,
            },
        )

    def send_global_metric(self, exp: str, round: int, metric: str, node: str, value: float) -> None:
        self._post(
            "/node-metric/global",
            {"node_id": self._node_id(node), "experiment": exp, "round": round, "metric": metric, "value": value},
        )

    def send_system_metric(self, node: str, metric: str, value: float, time: datetime.datetime) -> None:
        self._post(
            "/node-metric/system",
            {
                "node_id": self._node_id(node),
                "metric": metric,
                "value": value,
                "time": time.strftime("%Y-%m-%d %H:%M:%S"),
            },
        )

    def get_pending_actions(self) -> Optional[list]:
        """Not provided by the dashboard API (reference: p2pfl_web_services.py:267-269)."""
        return None
