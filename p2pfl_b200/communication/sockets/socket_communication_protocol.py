"""Framed-socket transport: the control plane between nodes of this framework on one host
(or a trusted network), without gRPC.

Why: with the train step on the device a federated round of the benchmark is under 3 ms, and a
Python gRPC unary call costs both ends 0.15-0.3 ms of interpreter time - seven peers' worth of
bookkeeping messages per round was more GIL time than the round's own device time.  This transport
carries the same four calls as the reference's service (``handshake``, ``disconnect``,
``send_message``, ``send_weights``; p2pfl/communication/grpc/proto/node.proto:50-57) and the same
protobuf bodies, over one persistent stream socket per neighbour:

    request   u32 length | u8 kind | body          kind 1 handshake, 2 disconnect (body = address, utf-8)
    response  u32 length | error text (utf-8)           3 message (node.Message), 4 weights (node.Weights)

A call is a write followed by a read on the caller's side; ``deliver_many`` writes to every target
first and collects the answers afterwards, so the round trips of a broadcast overlap.  The server
side is one accept thread plus one reader thread per inbound connection (blocking socket calls
release the GIL).  It is NOT wire-compatible with a reference node - those are reached through the
gRPC transport, which stays the default.

Addresses: ``host:port``, bare ``host`` (a free port is picked) or ``unix:///abs/path``.
"""

from __future__ import annotations

import os
import socket
import struct
import threading
from typing import Any, List, Optional, Tuple

from p2pfl_b200.communication.grpc import proto
from p2pfl_b200.communication.grpc.address import AddressParser
from p2pfl_b200.communication.grpc.grpc_communication_protocol import message_from_pb, to_pb, weights_from_pb
from p2pfl_b200.communication.messages import Message
from p2pfl_b200.communication.protocol import BaseProtocol, Envelope, Transport
from p2pfl_b200.exceptions import CommunicationError
from p2pfl_b200.settings import Settings
from p2pfl_b200.utils.lockcheck import make_condition, make_lock, make_rlock

K_HANDSHAKE, K_DISCONNECT, K_MESSAGE, K_WEIGHTS = 1, 2, 3, 4
_HDR = struct.Struct("<IB")
_LEN = struct.Struct("<I")


def _recv_exact(sock: socket.socket, n: int) -> bytes:
    buf = bytearray(n)
    view, got = memoryview(buf), 0
    while got < n:
        k = sock.recv_into(view[got:], n - got)
        if k == 0:
            raise ConnectionError("peer closed the connection")
        got += k
    return bytes(buf)


def _connect(addr: str, timeout: float) -> socket.socket:
    if addr.startswith("unix://"):
        s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        s.settimeout(timeout)
        s.connect(addr[len("unix://") :])
        return s
    p = AddressParser(addr)
    s = socket.create_connection((p.host, p.port), timeout=timeout)
    s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
    return s


class _Link:
    """One outbound connection; a lock keeps request / response pairs of different threads apart."""

    def __init__(self, addr: str, timeout: float) -> None:
        self.addr = addr
        self.sock = _connect(addr, timeout)
        self.lock = make_lock("socket.link")

    def write(self, kind: int, body: bytes) -> None:
        self.sock.sendall(_HDR.pack(len(body), kind) + body)

    def read(self) -> Optional[str]:
        (n,) = _LEN.unpack(_recv_exact(self.sock, 4))
        return _recv_exact(self.sock, n).decode() if n else None

    def call(self, kind: int, body: bytes, timeout: float) -> Optional[str]:
        with self.lock:
            self.sock.settimeout(timeout)
            self.write(kind, body)
            return self.read()

    def close(self) -> None:
        try:
            self.sock.close()
        except Exception:
            pass


class _LocalLink:
    """Link to a node of THIS process: delivery is a call into its engine on the caller's thread, as
    in the in-memory transport - no frame, no syscalls, no second thread.  With several simulated nodes
    per process (eight per GPU in config 4) seven of every eight hops of a line are of this kind."""

    def __init__(self, addr: str, peer: "SocketTransport") -> None:
        self.addr, self.peer = addr, peer

    def deliver(self, env: Envelope) -> Optional[str]:
        if self.peer._stopped.is_set() or _LOCAL.get(self.addr) is not self.peer:
            raise ConnectionError(f"{self.addr} is down")
        eng = self.peer.engine
        return eng.on_message(env) if isinstance(env, Message) else eng.on_weights(env)

    def close(self) -> None:
        pass


_LOCAL: dict = {}  # address -> SocketTransport of a running node of this process
_LOCAL_LOCK = make_lock("socket.local")


class SocketTransport(Transport):
    def __init__(self, addr: str = "127.0.0.1") -> None:
        self._addr = AddressParser(addr).get_parsed_address()
        self._listener: Optional[socket.socket] = None
        self._conns: List[socket.socket] = []
        self._lock = make_lock("socket.conns")
        self._stopped = threading.Event()
        self._stopped.set()

    def address(self) -> str:
        return self._addr

    # ------------------------------------------------------------------ server
    def start(self) -> None:
        try:
            if self._addr.startswith("unix://"):
                path = self._addr[len("unix://") :]
                if os.path.exists(path):
                    os.unlink(path)
                ls = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                ls.bind(path)
            else:
                p = AddressParser(self._addr)
                ls = socket.socket(socket.AF_INET6 if ":" in (p.host or "") else socket.AF_INET, socket.SOCK_STREAM)
                ls.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
                ls.bind((p.host, p.port))
            ls.listen(256)
        except OSError as e:
            raise CommunicationError(f"cannot bind {self._addr}: {e}") from e
        self._listener = ls
        self._stopped.clear()
        with _LOCAL_LOCK:
            _LOCAL[self._addr] = self
        threading.Thread(target=self._accept_loop, args=(ls,), name=f"socket-accept-{self._addr}", daemon=True).start()

    def _accept_loop(self, ls: socket.socket) -> None:
        while not self._stopped.is_set():
            try:
                conn, _ = ls.accept()
            except OSError:
                return
            if conn.family != socket.AF_UNIX:
                conn.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            with self._lock:
                self._conns.append(conn)
            threading.Thread(target=self._serve, args=(conn,), name=f"socket-conn-{self._addr}", daemon=True).start()

    def _serve(self, conn: socket.socket) -> None:
        eng = self.engine
        try:
            while not self._stopped.is_set():
                n, kind = _HDR.unpack(_recv_exact(conn, _HDR.size))
                body = _recv_exact(conn, n) if n else b""
                if kind == K_MESSAGE:
                    err = eng.on_message(message_from_pb(proto.PbMessage.FromString(body)))
                elif kind == K_WEIGHTS:
                    err = eng.on_weights(weights_from_pb(proto.PbWeights.FromString(body)))
                elif kind == K_HANDSHAKE:
                    err = eng.on_handshake(body.decode())
                elif kind == K_DISCONNECT:
                    eng.on_disconnect(body.decode())
                    err = None
                else:
                    err = f"unknown frame kind {kind}"
                out = err.encode() if err else b""
                conn.sendall(_LEN.pack(len(out)) + out)
        except (ConnectionError, OSError, struct.error):
            pass  # the peer went away (or this endpoint is stopping)
        finally:
            try:
                conn.close()
            except Exception:
                pass

    def stop(self) -> None:
        self._stopped.set()
        with _LOCAL_LOCK:
            if _LOCAL.get(self._addr) is self:
                del _LOCAL[self._addr]
        ls, self._listener = self._listener, None
        if ls is not None:
            try:
                ls.shutdown(socket.SHUT_RDWR)  # wakes the accept loop
            except Exception:
                pass
            try:
                ls.close()
            except Exception:
                pass
            if self._addr.startswith("unix://"):
                try:
                    os.unlink(self._addr[len("unix://") :])
                except OSError:
                    pass
        with self._lock:
            conns, self._conns = self._conns, []
        for c in conns:
            try:
                c.shutdown(socket.SHUT_RDWR)
            except Exception:
                pass
            try:
                c.close()
            except Exception:
                pass

    def wait_for_termination(self) -> None:
        self._stopped.wait()

    # ------------------------------------------------------------------ client
    def open_link(self, addr: str, handshake: bool) -> Any:
        peer = _LOCAL.get(addr) if Settings.SOCKET_LOCAL_SHORTCUT else None
        if peer is not None and peer is not self and not peer._stopped.is_set():
            if handshake:
                err = peer.engine.on_handshake(self._addr)
                if err:
                    raise CommunicationError(f"{addr} refused the handshake: {err}")
            return _LocalLink(addr, peer)
        try:
            link = _Link(addr, Settings.GRPC_TIMEOUT)
        except OSError as e:
            raise CommunicationError(f"cannot reach {addr}: {e}") from e
        if handshake:
            try:
                err = link.call(K_HANDSHAKE, self._addr.encode(), Settings.GRPC_TIMEOUT)
            except Exception as e:
                link.close()
                raise CommunicationError(f"handshake with {addr} failed: {e}") from e
            if err:
                link.close()
                raise CommunicationError(f"{addr} refused the handshake: {err}")
        return link

    def close_link(self, addr: str, link: Any, disconnect_msg: bool) -> None:
        if isinstance(link, _LocalLink):
            if disconnect_msg and not link.peer._stopped.is_set():
                try:
                    link.peer.engine.on_disconnect(self._addr)
                except Exception:
                    pass
            return
        try:
            if disconnect_msg:
                link.call(K_DISCONNECT, self._addr.encode(), Settings.GRPC_TIMEOUT)
        except Exception:
            pass
        finally:
            link.close()

    @staticmethod
    def _frame(env: Envelope) -> Tuple[int, bytes, float]:
        body = to_pb(env).SerializeToString()
        if isinstance(env, Message):
            return K_MESSAGE, body, Settings.GRPC_TIMEOUT
        return K_WEIGHTS, body, Settings.GRPC_TIMEOUT + len(body) / 50e6

    def deliver(self, addr: str, link: Any, env: Envelope) -> Optional[str]:
        if isinstance(link, _LocalLink):
            return link.deliver(env)
        kind, body, timeout = self._frame(env)
        return link.call(kind, body, timeout)

    def deliver_many(self, targets: List[Any], env: Envelope) -> List[Any]:
        """Write the frame to every target, then collect the answers: the round trips overlap.
        Links are locked in address order, so two broadcasts over the same links cannot deadlock."""
        out: List[Any] = [None] * len(targets)
        remote = [i for i in range(len(targets)) if not isinstance(targets[i][1], _LocalLink)]
        if remote:
            kind, body, timeout = self._frame(env)
            frame = _HDR.pack(len(body), kind) + body
        order = sorted(remote, key=lambda i: targets[i][0])
        held = []
        for i in order:
            link = targets[i][1]
            link.lock.acquire()
            try:
                link.sock.settimeout(timeout)
                link.sock.sendall(frame)
                held.append(i)
            except Exception as e:
                link.lock.release()
                out[i] = e
        # neighbours of this process are served while the remote answers travel
        for i in range(len(targets)):
            if isinstance(targets[i][1], _LocalLink):
                try:
                    out[i] = targets[i][1].deliver(env)
                except Exception as e:
                    out[i] = e
        for i in held:
            link = targets[i][1]
            try:
                out[i] = link.read()
            except Exception as e:
                out[i] = e
            finally:
                link.lock.release()
        return out


class SocketCommunicationProtocol(BaseProtocol):
    """``Node(model, data, protocol=SocketCommunicationProtocol)``."""

    def __init__(self, addr: str = "127.0.0.1") -> None:
        super().__init__(SocketTransport(addr))
