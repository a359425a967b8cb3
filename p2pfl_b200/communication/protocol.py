"""Communication protocol: interface and the transport-independent engine.

``CommunicationProtocol`` is the abstract surface the node and the stages use
(reference: p2pfl/communication/communication_protocol.py:27-190).

``BaseProtocol`` implements all of it once - neighbour table, heartbeats,
flooding gossip with TTL and duplicate suppression, command dispatch, eviction
on failure - on top of a small :class:`Transport` (open/close a link, deliver
one envelope, run a server).  The reference writes this logic twice, once per
transport (grpc_server.py:102-197 and memory_server.py:100-186); here the
in-memory, gRPC and NVLink protocols differ only in their ``Transport``.
"""

from __future__ import annotations

import abc
import threading
from typing import Any, Callable, Dict, List, Optional, Union

from p2pfl_b200.commands.command import Command
from p2pfl_b200.communication.gossiper import Gossiper
from p2pfl_b200.communication.heartbeater import HEARTBEAT_CMD, Heartbeater
from p2pfl_b200.communication.messages import Message, Weights, new_message_id
from p2pfl_b200.communication.neighbors import Neighbors
from p2pfl_b200.exceptions import CommunicationError, NeighborNotConnectedError
from p2pfl_b200.management import tracer
from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings
from p2pfl_b200.utils.lockcheck import make_lock

Envelope = Union[Message, Weights]


class CommunicationProtocol(abc.ABC):
    """What a node needs from its network stack."""

    @abc.abstractmethod
    def start(self) -> None: ...

    @abc.abstractmethod
    def stop(self) -> None: ...

    @abc.abstractmethod
    def add_command(self, cmds: Union[Command, List[Command]]) -> None: ...

    @abc.abstractmethod
    def build_msg(self, cmd: str, args: Optional[List[str]] = None, round: Optional[int] = None) -> Message: ...

    @abc.abstractmethod
    def build_weights(
        self, cmd: str, round: int, serialized_model: Any, contributors: Optional[List[str]] = None, weight: int = 1
    ) -> Weights: ...

    @abc.abstractmethod
    def send(self, nei: str, msg: Envelope, create_connection: bool = False) -> None: ...

    @abc.abstractmethod
    def broadcast(self, msg: Message, node_list: Optional[List[str]] = None) -> None: ...

    @abc.abstractmethod
    def connect(self, addr: str, non_direct: bool = False) -> bool: ...

    @abc.abstractmethod
    def disconnect(self, nei: str, disconnect_msg: bool = True) -> None: ...

    @abc.abstractmethod
    def get_neighbors(self, only_direct: bool = False) -> Dict[str, Any]: ...

    @abc.abstractmethod
    def get_address(self) -> str: ...

    @abc.abstractmethod
    def wait_for_termination(self) -> None: ...

    @abc.abstractmethod
    def gossip_weights(
        self,
        early_stopping_fn: Callable[[], bool],
        get_candidates_fn: Callable[[], List[str]],
        status_fn: Callable[[], Any],
        model_fn: Callable[[str], Any],
        period: Optional[float] = None,
        create_connection: bool = False,
        wait_fn: Optional[Callable[[float], None]] = None,
    ) -> None: ...


class Transport(abc.ABC):
    """L1: moves one envelope to one peer and runs the inbound endpoint.

    The engine installs itself with :meth:`attach`; inbound traffic is handed
    to ``engine.on_message / on_weights / on_handshake / on_disconnect``."""

    engine: "BaseProtocol"

    def attach(self, engine: "BaseProtocol") -> None:
        self.engine = engine

    @abc.abstractmethod
    def address(self) -> str:
        """Final address of this endpoint (port chosen, socket path fixed)."""

    @abc.abstractmethod
    def start(self) -> None: ...

    @abc.abstractmethod
    def stop(self) -> None: ...

    @abc.abstractmethod
    def wait_for_termination(self) -> None: ...

    @abc.abstractmethod
    def open_link(self, addr: str, handshake: bool) -> Any:
        """Open a link to ``addr``; raise on failure."""

    @abc.abstractmethod
    def close_link(self, addr: str, link: Any, disconnect_msg: bool) -> None: ...

    @abc.abstractmethod
    def deliver(self, addr: str, link: Any, env: Envelope) -> Optional[str]:
        """Send and wait for the answer; returns the peer's error string or
        None.  Raises on transport failure."""

    def deliver_many(self, targets: List[Any], env: Envelope) -> List[Any]:
        """Deliver one envelope to several ``(addr, link)`` targets; one result per target
        (error string, None, or the exception).  Transports that can overlap the
        round trips override this."""
        return _deliver_many_sequential(self, targets, env)


def _deliver_many_sequential(transport: "Transport", targets: List[Any], env: Envelope) -> List[Any]:
    out: List[Any] = []
    for addr, link in targets:
        try:
            out.append(transport.deliver(addr, link, env))
        except Exception as e:  # noqa: PERF203
            out.append(e)
    return out


def _thread_uses_cuda() -> None:
    try:
        from p2pfl_b200 import ops

        ops.thread_uses_cuda()
    except Exception:
        pass


class _ServerHandle:
    """``protocol._server.stop()`` is reached into by the fault-injection tests
    (reference: test/communication_test.py:195)."""

    def __init__(self, transport: Transport) -> None:
        self._t = transport

    def stop(self) -> None:
        self._t.stop()


class BaseProtocol(CommunicationProtocol):
    """Transport-independent protocol engine."""

    def __init__(self, transport: Transport) -> None:
        self._transport = transport
        transport.attach(self)
        self.addr = transport.address()
        self._commands: Dict[str, Command] = {}
        self._on_neighbors_change: Optional[Callable[[], None]] = None
        self._neighbors = Neighbors(
            self.addr, transport.open_link, transport.close_link, on_change=self._neighbors_changed
        )
        self._gossiper = Gossiper(self.addr, self.send)
        self._gossiper.bundle_fn = self._broadcast_bundle
        self._gossiper.bundle_to_fn = self._send_bundle_to
        self._heartbeater = Heartbeater(self.addr, self._neighbors, self._broadcast_beat)
        self._server = _ServerHandle(transport)
        from p2pfl_b200.commands.control import HeartbeatCommand, NeighborsHintCommand

        self._extended: set = set()  # peers known to understand this framework's extra commands
        self.add_command(HeartbeatCommand(self._heartbeater, self._note_extended))
        self.add_command(NeighborsHintCommand(self.note_direct_neighbors_of))
        self._running = False
        self._terminated = threading.Event()
        self._direct_of: Dict[str, Any] = {}  # node -> set of its direct neighbours (from its hints)
        self._direct_of_version: Dict[str, int] = {}
        self._hint_lock = make_lock("protocol.hint")
        self._hint_version = 0
        self._temp_links: Dict[str, Any] = {}  # cached links to non-neighbours (create_connection sends)

    # ------------------------------------------------------------ lifecycle
    def start(self) -> None:
        self._transport.start()
        self._heartbeater.start()
        self._gossiper.start()
        self._running = True

    def stop(self) -> None:
        self._running = False
        self._heartbeater.stop()
        self._gossiper.stop()
        self._neighbors.clear_neighbors()
        for addr, link in list(self._temp_links.items()):
            try:
                self._transport.close_link(addr, link, False)
            except Exception:
                pass
        self._temp_links.clear()
        self._transport.stop()
        self._terminated.set()

    def wait_for_termination(self) -> None:
        self._transport.wait_for_termination()

    def get_address(self) -> str:
        return self.addr

    def set_neighbors_listener(self, fn: Optional[Callable[[], None]]) -> None:
        """Called (from arbitrary threads) whenever the neighbour set changes."""
        self._on_neighbors_change = fn

    def _neighbors_changed(self) -> None:
        known = self._neighbors.get_all()
        with self._hint_lock:  # a node that left may come back as a new process: forget its hints
            for gone in [a for a in self._direct_of if a not in known]:
                self._direct_of.pop(gone, None)
                self._direct_of_version.pop(gone, None)
        self._advertise_direct()
        fn = self._on_neighbors_change
        if fn is not None:
            try:
                fn()
            except Exception:
                pass

    # ------------------------------------------------------------- commands
    def add_command(self, cmds: Union[Command, List[Command]]) -> None:
        for c in cmds if isinstance(cmds, list) else [cmds]:
            self._commands[c.get_name()] = c

    # -------------------------------------------------------------- builders
    def build_msg(self, cmd: str, args: Optional[List[Any]] = None, round: Optional[int] = None) -> Message:
        return Message(
            source=self.addr,
            cmd=cmd,
            args=[str(a) for a in (args or [])],
            round=-1 if round is None else int(round),
            ttl=Settings.TTL,
            hash=new_message_id(),
        )

    def build_weights(
        self, cmd: str, round: int, serialized_model: Any, contributors: Optional[List[str]] = None, weight: int = 1
    ) -> Weights:
        return Weights(
            source=self.addr,
            cmd=cmd,
            round=int(round),
            weights=serialized_model,
            contributors=list(contributors or []),
            weight=int(weight),
        )

    # ------------------------------------------------------------- outbound
    def send(self, nei: str, msg: Envelope, create_connection: bool = False) -> None:
        """Deliver to one neighbour.  Any failure or error answer evicts the
        neighbour (reference: grpc_client.py:159-179, Q11)."""
        temp_link = None
        try:
            try:
                entry = self._neighbors.get(nei)
            except KeyError:
                entry = None
            link = entry.link if entry is not None and entry.direct else None
            if link is None:
                if not create_connection:
                    raise NeighborNotConnectedError(f"{nei} is not a direct neighbour of {self.addr}")
                # links to non-neighbours (train-set peers) are kept for reuse instead of
                # being opened and torn down around every model send
                link = self._temp_links.get(nei)
                if link is None:
                    link = self._temp_links[nei] = self._transport.open_link(nei, False)
            with tracer.span("send " + msg.cmd, self.addr, "send", to=nei):
                err = self._transport.deliver(nei, link, msg)
            if err:
                raise CommunicationError(f"{nei} answered with error: {err}")
        except NeighborNotConnectedError:
            raise
        except Exception as e:
            logger.info(self.addr, f"send of '{msg.cmd}' to {nei} failed ({e}); removing neighbour")
            stale = self._temp_links.pop(nei, None)
            if stale is not None:
                try:
                    self._transport.close_link(nei, stale, False)
                except Exception:
                    pass
            self._neighbors.remove(nei, disconnect_msg=True)
        finally:
            if temp_link is not None:
                try:
                    self._transport.close_link(nei, temp_link, False)
                except Exception:
                    pass

    def broadcast(self, msg: Message, node_list: Optional[List[str]] = None) -> None:
        """Send to every direct neighbour (or the direct ones among ``node_list``); the
        round trips overlap when the transport supports it."""
        from p2pfl_b200.commands.control import REFERENCE_COMMANDS

        direct = self._neighbors.get_all(only_direct=True)
        names = [n for n in (node_list if node_list is not None else direct) if n in direct]
        if msg.cmd not in REFERENCE_COMMANDS:
            names = [n for n in names if n in self._extended]  # a reference node evicts on unknown commands
        targets = [(n, direct[n].link) for n in names]
        if not targets:
            return
        with tracer.span("broadcast " + msg.cmd, self.addr, "send", targets=len(targets)):
            results = self._transport.deliver_many(targets, msg)
        for (nei, _), res in zip(targets, results):
            if res is None:
                continue
            logger.info(self.addr, f"send of '{msg.cmd}' to {nei} failed ({res}); removing neighbour")
            self._neighbors.remove(nei, disconnect_msg=True)

    def broadcast_async(self, msg: Message, lazy: bool = False) -> None:
        """Fire-and-forget broadcast: queued to the gossiper thread, which performs the
        (overlapped) sends and evicts neighbours that fail, exactly like :meth:`broadcast`.
        ``lazy`` broadcasts may wait ``Settings.BUNDLE_LINGER`` for the next ordinary one;
        broadcasts that leave together are bundled (:meth:`_broadcast_bundle`)."""
        if not Settings.ASYNC_BOOKKEEPING:
            self.broadcast(msg)
            return
        targets = list(self._neighbors.get_all(only_direct=True))
        if targets:
            self._gossiper.add_broadcast(msg, self.broadcast, lazy=lazy)

    def multicast_async(self, msg: Message, names: List[str]) -> None:
        """Fire-and-forget delivery of ``msg`` to exactly ``names`` (direct neighbours or not), point to
        point: ttl 1, no flooding.  Links to non-neighbours are the cached temporary ones."""
        msg.ttl = 1
        targets = [n for n in names if n != self.addr]
        if not targets:
            return

        def deliver(m: Message) -> None:
            # all round trips overlap (one after the other, 15 trainers cost 15 RPC latencies per notice)
            pairs = []
            direct = self._neighbors.get_all(only_direct=True)
            for n in targets:
                try:
                    link = direct[n].link if n in direct else self._temp_links.get(n)
                    if link is None:
                        link = self._temp_links[n] = self._transport.open_link(n, False)
                    pairs.append((n, link))
                except Exception as e:  # one unreachable peer must not cost the others their copy
                    logger.debug(self.addr, f"multicast of '{m.cmd}': no link to {n}: {e}")
            if not pairs:
                return
            for (n, _), res in zip(pairs, self._transport.deliver_many(pairs, m)):
                if res is None:
                    continue
                logger.debug(self.addr, f"multicast of '{m.cmd}' to {n} failed: {res}")
                stale = self._temp_links.pop(n, None)
                if stale is not None:
                    try:
                        self._transport.close_link(n, stale, False)
                    except Exception:
                        pass

        if Settings.ASYNC_BOOKKEEPING:
            self._gossiper.add_broadcast(msg, deliver)
        else:
            deliver(msg)

    def _send_bundle_to(self, nei: str, msgs: List[Message]) -> None:
        """Several (relayed) messages for ONE neighbour: a single ``bundle`` if it is extended."""
        from p2pfl_b200.commands.control import BUNDLE_CMD, pack_bundled

        if nei in self._extended and len(msgs) > 1:
            bundle = self.build_msg(BUNDLE_CMD, [pack_bundled(m) for m in msgs])
            bundle.ttl = 1
            self.send(nei, bundle)
        else:
            for m in msgs:
                self.send(nei, m)

    def _broadcast_bundle(self, msgs: List[Message]) -> None:
        """Several broadcasts at once: extended neighbours get ONE ``bundle`` message (ttl 1)
        whose arguments are the packed messages - each is then handled, relayed and
        de-duplicated by the receiver as if it had arrived alone; reference neighbours get the
        messages one by one (they would evict us for a command they do not know)."""
        from p2pfl_b200.commands.control import BUNDLE_CMD, REFERENCE_COMMANDS, pack_bundled

        direct = self._neighbors.get_all(only_direct=True)
        ext = [(n, direct[n].link) for n in direct if n in self._extended]
        plain = [n for n in direct if n not in self._extended]
        failed: Dict[str, str] = {}
        if ext:
            bundle = self.build_msg(BUNDLE_CMD, [pack_bundled(m) for m in msgs])
            bundle.ttl = 1
            with tracer.span("broadcast bundle", self.addr, "send", targets=len(ext), messages=len(msgs)):
                results = self._transport.deliver_many(ext, bundle)
            for (nei, _), res in zip(ext, results):
                if res is not None:
                    failed[nei] = res
        if plain:
            for m in msgs:
                if m.cmd in REFERENCE_COMMANDS:
                    self.broadcast(m, node_list=plain)
        for nei, res in failed.items():
            logger.info(self.addr, f"send of a bundle to {nei} failed ({res}); removing neighbour")
            self._neighbors.remove(nei, disconnect_msg=True)

    def _beat_msg(self, now: float) -> Message:
        from p2pfl_b200.commands.control import EXTENDED_MARK

        return self.build_msg(HEARTBEAT_CMD, [str(now)], round=EXTENDED_MARK)

    def _broadcast_beat(self, now: float) -> None:
        self.broadcast(self._beat_msg(now))
        hint = self._hint_msg()
        if hint is not None:
            self.broadcast(hint)

    def _hint_msg(self) -> Optional[Message]:
        """``neighbors_hint``: tells extended direct neighbours whom I reach directly, so that they
        need not relay my broadcasts there (point-to-point, ttl 1: a reference node never sees it)."""
        with self._hint_lock:
            # version and list are taken together: of two hints built by two threads the one with the
            # higher version never carries the older list, whatever order they leave in
            self._hint_version += 1
            version = self._hint_version
            direct = list(self._neighbors.get_all(only_direct=True))
        if not Settings.RELAY_PRUNING or len(direct) > Settings.RELAY_PRUNING_MAX_LIST:
            return None
        hint = self.build_msg("neighbors_hint", sorted(direct), round=version)
        hint.ttl = 1
        return hint

    def _note_extended(self, addr: str) -> None:
        """``addr`` sent a beat carrying this framework's mark."""
        if addr in self._extended:
            return
        self._extended.add(addr)
        self._advertise_direct()

    def _advertise_direct(self) -> None:
        """The set of direct neighbours, or of those known to understand the hint, changed: advertise
        it NOW (queued to the gossiper thread).  Waiting for the next periodic beat left a freshly
        connected full mesh of 8 nodes flooding every broadcast through every other node for up to
        ``HEARTBEAT_PERIOD``: 7 copies of each message per receiver, 5 ms per round instead of 0.5
        (bench/control_plane_bench.py)."""
        if not self._running or not self._extended:
            return
        hint = self._hint_msg()
        if hint is not None:
            self._gossiper.add_broadcast(hint, self.broadcast)

    def _greet(self, addr: str) -> None:
        """Right after a link comes up: an immediate beat, so both ends learn at once whether
        the other speaks the extended commands (queued: never re-enter the peer synchronously)."""
        import time as _time

        self._gossiper.add_message(self._beat_msg(_time.time()), [addr])

    def note_direct_neighbors_of(self, source: str, names: List[str], version: int = -1) -> None:
        with self._hint_lock:
            if version >= 0 and version < self._direct_of_version.get(source, -1):
                return  # an older hint that was overtaken on its way
            self._direct_of_version[source] = version
            self._direct_of[source] = set(names)

    # ------------------------------------------------------------ neighbours
    def connect(self, addr: str, non_direct: bool = False) -> bool:
        ok = self._neighbors.add(addr, non_direct=non_direct)
        if ok and not non_direct and self._running:
            self._greet(addr)
        return ok

    def disconnect(self, nei: str, disconnect_msg: bool = True) -> None:
        self._neighbors.remove(nei, disconnect_msg=disconnect_msg)

    def get_neighbors(self, only_direct: bool = False) -> Dict[str, Any]:
        return self._neighbors.get_all(only_direct)

    def gossip_weights(self, *args: Any, **kwargs: Any) -> None:
        self._gossiper.gossip_weights(*args, **kwargs)

    # -------------------------------------------------------------- inbound
    def on_handshake(self, addr: str) -> Optional[str]:
        """Peer opened a link to us: add it as a direct neighbour without
        hand-shaking back (reference: grpc_server.py:102-114)."""
        if not self._running:
            return "node not running"
        if self._neighbors.add(addr, non_direct=False, handshake=False):
            self._greet(addr)
            return None
        # already a direct neighbour is fine (both sides connected at once)
        try:
            if self._neighbors.get(addr).direct:
                return None
        except KeyError:
            pass
        return "cannot add the neighbour"

    def on_disconnect(self, addr: str) -> None:
        self._neighbors.remove(addr, disconnect_msg=False)

    def on_message(self, msg: Message, via: Optional[str] = None) -> Optional[str]:
        """Flooding gossip: drop duplicates, relay while TTL lasts, dispatch
        (reference: grpc_server.py:130-166).  ``via``: the neighbour that handed us the message
        when that is known (bundled relays) - it is not sent back there."""
        if not self._running:
            return "node not running"
        if not self._gossiper.check_and_set_processed(msg.hash):
            return None
        if msg.cmd == "bundle":
            from p2pfl_b200.commands.control import unpack_bundled

            err = None
            for packed in msg.args:
                try:
                    inner = unpack_bundled(msg.source, packed)
                    if inner.cmd == "bundle":  # bundles do not nest
                        err = "nested bundle"
                        continue
                    err = self.on_message(inner, via=msg.source) or err
                except Exception as e:
                    err = f"bad bundled message: {e}"
            return err
        if msg.ttl > 1:
            # flood to my direct neighbours, except the originator and - topology-aware
            # pruning - those the originator advertises it reaches directly
            from p2pfl_b200.commands.control import REFERENCE_COMMANDS

            covered = self._direct_of.get(msg.source, ()) if Settings.RELAY_PRUNING else ()
            targets = [
                n for n in self._neighbors.get_all(only_direct=True) if n != msg.source and n != via and n not in covered
            ]
            if msg.cmd not in REFERENCE_COMMANDS:
                targets = [n for n in targets if n in self._extended]
            if targets:
                self._gossiper.add_message(msg.relay_copy(), targets)
        cmd = self._commands.get(msg.cmd)
        if cmd is None:
            logger.error(self.addr, f"unknown command '{msg.cmd}' from {msg.source}")
            return f"Unknown command: {msg.cmd}"
        try:
            with tracer.span(msg.cmd, self.addr, "message", source=msg.source, round=msg.round):
                cmd.execute(msg.source, msg.round, *msg.args)
        except Exception as e:
            logger.error(self.addr, f"command '{msg.cmd}' from {msg.source} failed: {e!r}")
            return f"Error while processing command: {msg.cmd} {msg.args}: {e}"
        return None

    def on_weights(self, w: Weights) -> Optional[str]:
        """Model message: no dedupe, no relay (reference: grpc_server.py:168-197)."""
        if not self._running:
            return "node not running"
        _thread_uses_cuda()  # the handler may pull / decode a model on the device
        cmd = self._commands.get(w.cmd)
        if cmd is None:
            return f"Unknown command: {w.cmd}"
        try:
            with tracer.span(w.cmd, self.addr, "model", source=w.source, round=w.round, contributors=len(w.contributors)):
                cmd.execute(w.source, w.round, weights=w.weights, contributors=list(w.contributors), weight=w.weight)
        except Exception as e:
            logger.error(self.addr, f"weights command '{w.cmd}' from {w.source} failed: {e!r}")
            return f"Error while processing model: {w.cmd}: {e}"
        return None
