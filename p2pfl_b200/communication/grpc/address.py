"""Address parsing for the gRPC transport
(reference: p2pfl/communication/grpc/address.py:43-99).

Accepted forms: ``unix:///abs/path``, ``host:port``, ``[v6]:port``, bare
``host`` / ``[v6]`` (a free port is picked by binding port 0)."""

from __future__ import annotations

import socket
from typing import Optional, Tuple


class AddressParser:
    def __init__(self, address: str) -> None:
        self.unix_domain = False
        self.host: Optional[str] = None
        self.port: Optional[int] = None
        self._parsed = self._parse(address)

    def _parse(self, address: str) -> str:
        if address.startswith("unix://"):
            self.unix_domain = True
            return address
        host, port = self._split(address)
        if port is None:
            port = self._free_port(host)
        self.host, self.port = host, port
        if ":" in host:  # IPv6 literal
            return f"[{host}]:{port}"
        return f"{host}:{port}"

    @staticmethod
    def _split(address: str) -> Tuple[str, Optional[int]]:
        if address.startswith("["):
            end = address.find("]")
            if end < 0:
                raise ValueError(f"bad IPv6 address: {address}")
            host = address[1:end]
            rest = address[end + 1 :]
            if rest.startswith(":") and rest[1:].isdigit():
                return host, int(rest[1:])
            if rest:
                raise ValueError(f"bad address: {address}")
            return host, None
        if address.count(":") == 1:
            host, port = address.split(":")
            if not port.isdigit():
                raise ValueError(f"bad port in address: {address}")
            return host, int(port)
        if address.count(":") > 1:  # bare IPv6
            return address, None
        return address, None

    @staticmethod
    def _free_port(host: str) -> int:
        family = socket.AF_INET6 if ":" in host else socket.AF_INET
        with socket.socket(family, socket.SOCK_STREAM) as s:
            s.bind(("", 0))
            return int(s.getsockname()[1])

    def get_parsed_address(self) -> str:
        return self._parsed
