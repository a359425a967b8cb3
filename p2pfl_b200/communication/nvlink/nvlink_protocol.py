"""NVLink device data plane + host control plane.

The control plane (votes, heartbeats, ``models_aggregated`` ... all the small
messages, and the *envelope* of every model message) is the ordinary protocol
engine over gRPC (one process per GPU) or over the in-process transport (many
nodes in one process).  What changes is how a MODEL moves:

* ``fused_exchange``  - the product path.  When every member of the round's
  train set has published a device arena, each trainer launches ONE kernel
  (``p2pfl_gossip_fused``) that stores its model tile by tile into slot[me] of
  every other trainer's inbox through peer-mapped pointers (NVLink/NVSwitch),
  then accumulates the tiles the others stored into its own inbox, weighted by
  sample count, straight into the live parameters.  No NCCL, no host copy, no
  pickle (replaces gossiper.py:239 + fedavg.py:49-58 + lightning_learner.py:152).
* ``send`` of a ``DevicePayload`` - the general path (init weights, diffusion,
  partial aggregates, any topology).  The control message carries a ~200-byte
  ``DeviceRef``; the receiver *pulls* the tensor from the sender's arena with a
  kernel that reads peer memory over NVLink (in-process: a device-to-device
  read), so no inbox slot can be clobbered by two senders.
* anything else (a peer without an arena, a reference node) falls back to the
  pickled host payload, byte-compatible with the reference.

Arena discovery: a node announces ``arena_info`` (geometry, layout hash, GPU,
pid, CUDA IPC handle, sample count) with a TTL-flooded control message when its
learner is built; peers map the arena lazily (same process: the pointer itself;
other process: ``cudaIpcOpenMemHandle``).
"""

from __future__ import annotations

import os
import socket
import threading
import time
from typing import Any, Dict, List, Optional, Tuple

import torch

from p2pfl_b200 import ops
from p2pfl_b200.commands.command import Command
from p2pfl_b200.communication.grpc.grpc_communication_protocol import GrpcTransport
from p2pfl_b200.communication.memory.memory_communication_protocol import MemoryTransport
from p2pfl_b200.communication.messages import Weights
from p2pfl_b200.communication.protocol import BaseProtocol, Envelope, Transport
from p2pfl_b200.learning.params import FlatParams
from p2pfl_b200.management.logger import logger
from p2pfl_b200.parallel.arena import ArenaGeometry, DeviceArena, PeerArena, tensor_from_ptr
from p2pfl_b200.parallel.payload import DevicePayload, DeviceRef
from p2pfl_b200.settings import Settings

_HOST = socket.gethostname()
_PID = os.getpid()

# arenas of nodes living in THIS process: addr -> DeviceArena
_local_arenas: Dict[str, DeviceArena] = {}
_local_lock = threading.Lock()


class ArenaInfoCommand(Command):
    """``arena_info [geo_key, layout_hash, device, pid, host, ipc_hex, n_samples]``."""

    def __init__(self, proto: "NvlinkCommunicationProtocol") -> None:
        self._proto = proto

    @staticmethod
    def get_name() -> str:
        return "arena_info"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        self._proto._on_arena_info(source, list(args))


class PeerInfo:
    def __init__(self, args: List[str]) -> None:
        self.geo = ArenaGeometry.from_key(args[0])
        self.layout_hash = args[1]
        self.device = int(args[2])
        self.pid = int(args[3])
        self.host = args[4]
        self.ipc = bytes.fromhex(args[5])
        self.n_samples = int(args[6])
        self.mapped: Optional[PeerArena] = None


class NvlinkCommunicationProtocol(BaseProtocol):
    """``Node(model, data, protocol=NvlinkCommunicationProtocol, learner=B200Learner)``."""

    device_payloads = True
    control = "grpc"

    def __init__(self, addr: str = "127.0.0.1") -> None:
        transport: Transport = GrpcTransport(addr) if self.control == "grpc" else MemoryTransport(addr)
        super().__init__(transport)
        self._node: Any = None
        self._peers: Dict[str, PeerInfo] = {}
        self._peers_lock = threading.Lock()
        self._nonce = 0
        self._no_arena: set = set()
        self.add_command(ArenaInfoCommand(self))
        # counts, bytes and DEVICE time (CUDA events on the stream that did the work) per path
        self.stats = {"fused": 0, "pulled": 0, "host": 0, "fused_bytes": 0, "pulled_bytes": 0, "fused_ms": 0.0, "pulled_ms": 0.0}

    # ------------------------------------------------------------------ wiring
    def bind_node(self, node: Any) -> None:
        self._node = node

    def stop(self) -> None:
        super().stop()
        with self._peers_lock:
            peers, self._peers = list(self._peers.values()), {}
        for p in peers:
            if p.mapped is not None:
                try:
                    p.mapped.close()
                except Exception:
                    pass
        with _local_lock:
            _local_arenas.pop(self.addr, None)

    def _learner(self) -> Any:
        return None if self._node is None else self._node.state.learner

    # ---------------------------------------------------------------- discovery
    def publish_arena(self, learner: Any) -> None:
        """Announce this node's arena to everybody (called when the learner is built)."""
        arena = getattr(learner, "arena", None)
        if arena is None:
            return
        with _local_lock:
            _local_arenas[self.addr] = arena
        learner.plane = self
        try:
            ipc = arena.ipc_handle().hex()
        except Exception as e:  # IPC export can be refused in restricted containers
            logger.debug(self.addr, f"arena IPC export failed: {e}")
            ipc = ""
        args = [arena.geo.key(), learner.layout.hash, str(arena.device), str(_PID), _HOST, ipc, str(learner.get_num_samples()[0])]
        self._my_info = args
        self.broadcast(self.build_msg("arena_info", args))

    def _on_arena_info(self, source: str, args: List[str]) -> None:
        info = PeerInfo(args)
        with self._peers_lock:
            old = self._peers.get(source)
            if old is not None and old.ipc == info.ipc and old.pid == info.pid and old.geo == info.geo:
                old.n_samples, old.layout_hash = info.n_samples, info.layout_hash
                return  # same arena as before (pooled): keep the mapping
            self._peers[source] = info
            self._no_arena.discard(source)
        if old is not None and old.mapped is not None:
            old.mapped.close()
        if self._node is not None:
            self._node.state.notify()

    def _map_peer(self, addr: str) -> Optional[PeerArena]:
        """PeerArena for ``addr`` (None if it has none or it is unreachable from here)."""
        learner = self._learner()
        my_dev = getattr(getattr(learner, "arena", None), "device", None)
        if my_dev is None:
            return None
        with self._peers_lock:
            info = self._peers.get(addr)
            if info is None or info.host != _HOST:
                return None
            if info.mapped is not None:
                return info.mapped
            try:
                if info.pid == _PID:
                    with _local_lock:
                        arena = _local_arenas.get(addr)
                    if arena is None:
                        return None
                    if arena.device != my_dev:
                        ops.rt_enable_peer(my_dev, arena.device)
                    info.mapped = PeerArena(arena.base, arena.geo, arena.device)
                else:
                    if not info.ipc:
                        return None
                    info.mapped = PeerArena.open_ipc(info.ipc, info.geo, my_dev, info.device)
            except Exception as e:
                logger.warning(self.addr, f"cannot map the arena of {addr}: {e}")
                return None
            return info.mapped

    def peer_samples(self, addr: str) -> Optional[int]:
        with self._peers_lock:
            info = self._peers.get(addr)
            return None if info is None else info.n_samples

    # ------------------------------------------------------------------ outbound
    def send(self, nei: str, msg: Envelope, create_connection: bool = False) -> None:
        if isinstance(msg, Weights) and isinstance(msg.weights, DevicePayload):
            msg = self._lower_payload(nei, msg)
        super().send(nei, msg, create_connection)

    def _lower_payload(self, nei: str, w: Weights) -> Weights:
        """DevicePayload -> DeviceRef (peer can pull) or pickled bytes (it cannot)."""
        payload: DevicePayload = w.weights
        learner = payload.owner
        arena = getattr(learner, "arena", None)
        with self._peers_lock:
            info = self._peers.get(nei)
        if info is None and arena is not None and nei not in self._no_arena and self._node is not None:
            # the peer may still be building its learner: its announcement is milliseconds away
            self._node.state.wait_for(lambda: nei in self._peers, 0.5)
            with self._peers_lock:
                info = self._peers.get(nei)
            if info is None:
                self._no_arena.add(nei)
        reachable = (
            arena is not None
            and info is not None
            and info.host == _HOST
            and info.layout_hash == learner.layout.hash
            and (info.pid == _PID or bool(getattr(self, "_my_info", ["", "", "", "", "", ""])[5]))
        )
        if not reachable:
            self.stats["host"] += 1
            return Weights(w.source, w.cmd, w.round, learner.encode_parameters(params=payload.params), w.contributors, w.weight)
        flat = payload.params.flat
        if flat.data_ptr() == arena.params.data_ptr():
            region = "params"
        else:
            # stage into the exportable outbox (the sender keeps it stable until the RPC returns)
            with torch.cuda.stream(learner.stream):
                arena.outbox.copy_(flat.to(torch.float32), non_blocking=True)
            learner.stream.synchronize()
            region = "outbox"
        self._nonce += 1
        extra = {k: v.tolist() for k, v in payload.params.extra.items()}
        ref = DeviceRef(self.addr, region, learner.layout.hash, int(flat.numel()), extra, self._nonce)
        return Weights(w.source, w.cmd, w.round, ref.to_bytes(), w.contributors, w.weight)

    # ------------------------------------------------------------------- inbound
    # A ``DeviceRef`` travels through the ordinary command path untouched; the device
    # learner resolves it in ``decode_parameters`` by calling :meth:`pull`.  That way a
    # reference that arrives before the learner exists is stashed like any other payload.
    def pull(self, ref: DeviceRef) -> Any:
        return self._pull(ref)

    def _pull(self, ref: DeviceRef) -> Any:
        """Read the sender's tensor over NVLink into a private bf16/fp32 buffer."""
        learner = self._learner()
        if learner is None or getattr(learner, "arena", None) is None:
            raise RuntimeError("no device learner to receive into (learner not built yet)")
        if ref.layout_hash != learner.layout.hash:
            from p2pfl_b200.exceptions import ModelNotMatchingError

            raise ModelNotMatchingError("layout hash differs from the local model")
        peer = self._map_peer(ref.addr)
        if peer is None:
            raise RuntimeError(f"arena of {ref.addr} is not mapped")
        dev = learner.arena.device
        src = tensor_from_ptr(peer.region_ptr(ref.region), ref.n, torch.float32, dev)
        # pulled models stay fp32 (exact): this path carries initial weights, diffusion and
        # partial aggregates, none of which is bandwidth critical; the wire dtype applies
        # to the fused exchange
        comm = learner.comm_stream  # not the compute stream: that one may be capturing a graph
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(comm):
            out = torch.empty(ref.n, dtype=torch.float32, device=learner.device)
            e0.record(comm)
            ops.fedavg_into([src], [1.0], out, None)  # copy kernel reading peer memory over NVLink
            e1.record(comm)
        comm.synchronize()  # the sender may reuse its buffer once the RPC returns
        self.stats["pulled"] += 1
        self.stats["pulled_bytes"] += ref.n * 4
        self.stats["pulled_ms"] += e0.elapsed_time(e1)
        extra = {k: torch.tensor(v) for k, v in ref.extra.items()}
        return FlatParams(out, learner.layout, extra)

    # ------------------------------------------------------------ fused exchange
    def plan_fused_exchange(self, state: Any, learner: Any, wait: float = 2.0) -> Optional["FusedExchange"]:
        """A plan when every train-set member's arena is mapped, else None."""
        if not Settings.FUSED_EXCHANGE:
            return None
        arena = getattr(learner, "arena", None)
        train_set = sorted(state.train_set)
        if arena is None or state.addr not in train_set or len(train_set) < 2:
            return None
        others = [n for n in train_set if n != state.addr]
        if len(train_set) > arena.geo.slots or len(others) > ops.MAX_PEERS:
            return None
        deadline = time.time() + wait

        def known() -> bool:
            with self._peers_lock:
                return all(n in self._peers for n in others)

        state.wait_for(known, max(0.0, deadline - time.time()))
        peers: Dict[str, PeerArena] = {}
        samples: Dict[str, int] = {state.addr: learner.get_num_samples()[0]}
        for n in others:
            with self._peers_lock:
                info = self._peers.get(n)
            if info is None or info.layout_hash != learner.layout.hash or info.geo != arena.geo:
                return None
            if info.device == arena.device:
                # co-located on my GPU: kernels that wait on one another must not share a device
                # (B200_PROFILING.md); such train sets use the pull path
                return None
            p = self._map_peer(n)
            if p is None:
                return None
            peers[n] = p
            samples[n] = info.n_samples
        return FusedExchange(self, state, learner, train_set, peers, samples)


class FusedExchange:
    """One all-to-all push + FedAvg among the train set as a single kernel per node."""

    def __init__(self, proto: NvlinkCommunicationProtocol, state: Any, learner: Any, train_set: List[str],
                 peers: Dict[str, PeerArena], samples: Dict[str, int]) -> None:  # fmt: skip
        self.proto, self.state, self.learner = proto, state, learner
        self.train_set, self.peers, self.samples = train_set, peers, samples
        self.complete = False

    def run(self, seq: int) -> Tuple[List[str], int]:
        """Launch, wait, and return (contributors included, total weight)."""
        learner, arena = self.learner, self.learner.arena
        me = self.state.addr
        my_slot = self.train_set.index(me)
        bank = seq & 1
        others = [n for n in self.train_set if n != me]
        total = float(sum(self.samples[n] for n in self.train_set))
        wire_bf16 = arena.geo.wire_bf16
        push = ops.make_push_table(
            [self.peers[n].slot_ptr(my_slot, bank) for n in others],
            [self.peers[n].flag_ptr(my_slot, bank) for n in others],
            wire_bf16,
        )
        red = ops.make_reduce_table(
            [arena.slot_ptr(self.train_set.index(n), bank) for n in others],
            [arena.flag_ptr(self.train_set.index(n), bank) for n in others],
            [self.samples[n] / total for n in others],
            my_slot,  # summation in train-set order, local model at its own rank => identical bits on every node
            self.samples[me] / total,
            wire_bf16,
            True,
        )
        clock_hz = 1.9e9
        timeout_cycles = int(Settings.AGGREGATION_TIMEOUT * clock_hz)
        began = torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(learner.stream):
            arena.status.zero_()
            arena.abort.zero_()
            began.record(learner.stream)
            ops.gossip_fused(arena.params, learner.w16, arena.geo.tile, seq, push, red, arena.abort, timeout_cycles, arena.status)
            done = torch.cuda.Event(enable_timing=True)
            done.record(learner.stream)
        # host watchdog: flip the abort word if a contributor disappears or the experiment stops
        t0 = time.time()
        while not done.query():
            gone = [n for n in others if n not in self.proto.get_neighbors(only_direct=False)]
            if gone or self.state.round is None or time.time() - t0 > Settings.AGGREGATION_TIMEOUT:
                with torch.cuda.stream(self.proto._abort_stream(learner)):
                    arena.abort.fill_(1)
                done.synchronize()
                break
            time.sleep(0.0002)
        status = arena.status.cpu()
        learner._params_written_by_plane()
        self.proto.stats["fused"] += 1
        self.proto.stats["fused_ms"] += began.elapsed_time(done)  # includes the wait for the slowest peer's tiles
        self.proto.stats["fused_bytes"] += arena.n * (2 if wire_bf16 else 4) * len(others)
        self.complete = int(status[0]) == 0
        if not self.complete:
            logger.warning(me, "fused exchange finished with missing contributors (timeout/abort)")
        return list(self.train_set), int(total)


def _abort_stream(self: NvlinkCommunicationProtocol, learner: Any) -> torch.cuda.Stream:
    s = getattr(self, "_abort_s", None)
    if s is None:
        s = self._abort_s = torch.cuda.Stream(device=learner.device)
    return s


NvlinkCommunicationProtocol._abort_stream = _abort_stream  # type: ignore[attr-defined]


class NvlinkMemoryProtocol(NvlinkCommunicationProtocol):
    """Same data plane, in-process control plane (several nodes per process / per GPU)."""

    control = "memory"
