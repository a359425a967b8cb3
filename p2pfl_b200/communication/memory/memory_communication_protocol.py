"""In-process transport: every node of the process registers an endpoint in a
process-global registry and a send is a direct method call on the receiver's
engine, on the sender's thread (reference: p2pfl/communication/memory/*.py).

Fixes over the reference: stopping one endpoint removes only its own registry
key (Q7); relays copy the envelope (Q6, done in ``Message.relay_copy``).
"""

from __future__ import annotations

import itertools
import threading
from typing import Any, Dict, Optional

from p2pfl_b200.communication.protocol import BaseProtocol, Envelope, Transport
from p2pfl_b200.communication.messages import Message
from p2pfl_b200.exceptions import CommunicationError
from p2pfl_b200.utils.lockcheck import make_condition, make_lock, make_rlock


class ServerSingleton:
    """addr -> live endpoint (reference: memory/server_singleton.py:22-43)."""

    _endpoints: Dict[str, "MemoryTransport"] = {}
    _lock = make_lock("memory.registry")
    _ids = itertools.count()

    @classmethod
    def register(cls, addr: str, ep: "MemoryTransport") -> None:
        with cls._lock:
            if addr in cls._endpoints:
                raise ValueError(f"address {addr} already in use")
            cls._endpoints[addr] = ep

    @classmethod
    def unregister(cls, addr: str, ep: "MemoryTransport") -> None:
        with cls._lock:
            if cls._endpoints.get(addr) is ep:
                del cls._endpoints[addr]

    @classmethod
    def lookup(cls, addr: str) -> "MemoryTransport":
        with cls._lock:
            ep = cls._endpoints.get(addr)
        if ep is None:
            raise CommunicationError(f"no in-memory node at {addr}")
        return ep

    @classmethod
    def fresh_address(cls, prefix: str = "node") -> str:
        return f"{prefix}-{next(cls._ids)}"

    @classmethod
    def reset(cls) -> None:
        with cls._lock:
            cls._endpoints.clear()


class MemoryTransport(Transport):
    def __init__(self, addr: Optional[str] = None) -> None:
        # any string is a valid address; "127.0.0.1" (Node's default) means "pick one"
        if addr is None or addr == "127.0.0.1":
            addr = ServerSingleton.fresh_address()
        if " " in addr:
            raise ValueError("node addresses must not contain spaces")
        self._addr = addr
        self._up = False
        self._done = threading.Event()

    def address(self) -> str:
        return self._addr

    def start(self) -> None:
        ServerSingleton.register(self._addr, self)
        self._up = True
        self._done.clear()

    def stop(self) -> None:
        self._up = False
        ServerSingleton.unregister(self._addr, self)
        self._done.set()

    def wait_for_termination(self) -> None:
        self._done.wait()

    def _peer(self, addr: str) -> "MemoryTransport":
        ep = ServerSingleton.lookup(addr)
        if not ep._up:
            raise CommunicationError(f"{addr} is down")
        return ep

    def open_link(self, addr: str, handshake: bool) -> Any:
        ep = self._peer(addr)
        if handshake:
            err = ep.engine.on_handshake(self._addr)
            if err:
                raise CommunicationError(err)
        return addr  # the link is just the name; endpoints are looked up per send

    def close_link(self, addr: str, link: Any, disconnect_msg: bool) -> None:
        if disconnect_msg:
            try:
                self._peer(addr).engine.on_disconnect(self._addr)
            except CommunicationError:
                pass

    def deliver(self, addr: str, link: Any, env: Envelope) -> Optional[str]:
        ep = self._peer(addr)
        if isinstance(env, Message):
            return ep.engine.on_message(env)
        return ep.engine.on_weights(env)


class InMemoryCommunicationProtocol(BaseProtocol):
    """``Node(model, data, protocol=InMemoryCommunicationProtocol)``."""

    def __init__(self, addr: Optional[str] = None) -> None:
        super().__init__(MemoryTransport(addr))
