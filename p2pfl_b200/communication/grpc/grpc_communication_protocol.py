"""gRPC transport, wire-compatible with the reference
(p2pfl/communication/grpc/{grpc_client,grpc_server,grpc_neighbors}.py).

Four unary RPCs on service ``node.NodeServices``: ``handshake``,
``disconnect``, ``send_message``, ``send_weights``.  Differences from the
reference: send/receive size limits are raised on both channel and server so
models over 4 MiB fit (Q9); the server pool size is a setting.  All protocol
logic (dedupe, relay, dispatch, eviction) lives in ``BaseProtocol``; this file
only marshals envelopes.
"""

from __future__ import annotations

import threading
from concurrent import futures
from typing import Any, Dict, Optional

import grpc

from p2pfl_b200.communication.grpc import proto
from p2pfl_b200.communication.grpc.address import AddressParser
from p2pfl_b200.communication.messages import Message, Weights
from p2pfl_b200.communication.protocol import BaseProtocol, Envelope, Transport
from p2pfl_b200.exceptions import CommunicationError
from p2pfl_b200.settings import Settings


def _channel_options() -> list:
    n = Settings.GRPC_MAX_MESSAGE_BYTES
    return [("grpc.max_send_message_length", n), ("grpc.max_receive_message_length", n)]


def to_pb(env: Envelope) -> Any:
    if isinstance(env, Message):
        pb = proto.PbMessage(source=env.source, ttl=env.ttl, hash=env.hash, cmd=env.cmd, args=env.args)
        pb.round = env.round
        return pb
    if not isinstance(env.weights, (bytes, bytearray, memoryview)):
        raise CommunicationError("the gRPC transport carries host payloads (bytes) only")
    return proto.PbWeights(
        source=env.source,
        round=env.round,
        weights=bytes(env.weights),
        contributors=env.contributors,
        weight=env.weight,
        cmd=env.cmd,
    )


def message_from_pb(pb: Any) -> Message:
    return Message(
        source=pb.source,
        cmd=pb.cmd,
        args=list(pb.args),
        round=pb.round if pb.HasField("round") else -1,
        ttl=pb.ttl,
        hash=pb.hash,
    )


def weights_from_pb(pb: Any) -> Weights:
    return Weights(
        source=pb.source,
        cmd=pb.cmd,
        round=pb.round,
        weights=pb.weights,
        contributors=list(pb.contributors),
        weight=pb.weight,
    )


class _Link:
    """One open channel with its four method callables."""

    def __init__(self, addr: str) -> None:
        self.channel = grpc.insecure_channel(addr, options=_channel_options())
        self.calls: Dict[str, Any] = {}
        for name, (req, resp) in proto.METHODS.items():
            self.calls[name] = self.channel.unary_unary(
                proto.method_path(name),
                request_serializer=req.SerializeToString,
                response_deserializer=resp.FromString,
            )

    def close(self) -> None:
        self.channel.close()


class GrpcTransport(Transport):
    def __init__(self, addr: str = "127.0.0.1") -> None:
        self._addr = AddressParser(addr).get_parsed_address()
        self._server: Optional[grpc.Server] = None
        self._lock = threading.Lock()

    def address(self) -> str:
        return self._addr

    # ------------------------------------------------------------- server
    def _handlers(self) -> grpc.GenericRpcHandler:
        eng = self.engine

        def handshake(req: Any, ctx: Any) -> Any:
            err = eng.on_handshake(req.addr)
            return proto.PbResponseMessage(error=err) if err else proto.PbResponseMessage()

        def disconnect(req: Any, ctx: Any) -> Any:
            eng.on_disconnect(req.addr)
            return proto.PbEmpty()

        def send_message(req: Any, ctx: Any) -> Any:
            err = eng.on_message(message_from_pb(req))
            return proto.PbResponseMessage(error=err) if err else proto.PbResponseMessage()

        def send_weights(req: Any, ctx: Any) -> Any:
            err = eng.on_weights(weights_from_pb(req))
            return proto.PbResponseMessage(error=err) if err else proto.PbResponseMessage()

        fns = {"handshake": handshake, "disconnect": disconnect, "send_message": send_message, "send_weights": send_weights}
        table = {
            name: grpc.unary_unary_rpc_method_handler(
                fns[name],
                request_deserializer=proto.METHODS[name][0].FromString,
                response_serializer=proto.METHODS[name][1].SerializeToString,
            )
            for name in fns
        }
        return grpc.method_handlers_generic_handler(proto.SERVICE, table)

    def start(self) -> None:
        with self._lock:
            server = grpc.server(
                futures.ThreadPoolExecutor(max_workers=max(2, Settings.GRPC_SERVER_WORKERS)),
                options=_channel_options(),
            )
            server.add_generic_rpc_handlers((self._handlers(),))
            try:
                bound = server.add_insecure_port(self._addr)
            except Exception as e:
                raise CommunicationError(f"cannot bind {self._addr}: {e}") from e
            if not bound and not self._addr.startswith("unix:"):
                raise CommunicationError(f"cannot bind {self._addr}")
            server.start()
            self._server = server

    def stop(self) -> None:
        with self._lock:
            server, self._server = self._server, None
        if server is not None:
            server.stop(0)

    def wait_for_termination(self) -> None:
        server = self._server
        if server is not None:
            server.wait_for_termination()

    # ------------------------------------------------------------- client
    def open_link(self, addr: str, handshake: bool) -> Any:
        link = _Link(addr)
        if handshake:
            try:
                res = link.calls["handshake"](proto.PbHandShakeRequest(addr=self._addr), timeout=Settings.GRPC_TIMEOUT)
            except Exception as e:
                link.close()
                raise CommunicationError(f"handshake with {addr} failed: {e}") from e
            if res.HasField("error"):
                link.close()
                raise CommunicationError(f"{addr} refused the handshake: {res.error}")
        return link

    def close_link(self, addr: str, link: Any, disconnect_msg: bool) -> None:
        try:
            if disconnect_msg:
                link.calls["disconnect"](proto.PbHandShakeRequest(addr=self._addr), timeout=Settings.GRPC_TIMEOUT)
        except Exception:
            pass
        finally:
            link.close()

    def deliver(self, addr: str, link: Any, env: Envelope) -> Optional[str]:
        pb = to_pb(env)
        name = "send_message" if isinstance(env, Message) else "send_weights"
        # large models get a proportionally longer deadline than control messages
        timeout = Settings.GRPC_TIMEOUT
        if name == "send_weights":
            timeout += len(pb.weights) / 50e6
        res = link.calls[name](pb, timeout=timeout)
        return res.error if res.HasField("error") else None


    def _call(self, env: Envelope):
        pb = to_pb(env)
        name = "send_message" if isinstance(env, Message) else "send_weights"
        timeout = Settings.GRPC_TIMEOUT
        if name == "send_weights":
            timeout += len(pb.weights) / 50e6
        return name, pb, timeout

    def deliver_many(self, targets, env):  # type: ignore[override]
        """All RPCs are issued as futures first and collected afterwards."""
        name, pb, timeout = self._call(env)
        futs = []  # (future | None, issue error | None); grpc futures are themselves RpcError instances
        for addr, link in targets:
            try:
                futs.append((link.calls[name].future(pb, timeout=timeout), None))
            except Exception as e:
                futs.append((None, e))
        out = []
        for f, err in futs:
            if f is None:
                out.append(err)
                continue
            try:
                res = f.result()
                out.append(res.error if res.HasField("error") else None)
            except Exception as e:
                out.append(e)
        return out


class GrpcCommunicationProtocol(BaseProtocol):
    """``Node(model, data, protocol=GrpcCommunicationProtocol)`` (the default)."""

    def __init__(self, addr: str = "127.0.0.1") -> None:
        super().__init__(GrpcTransport(addr))
