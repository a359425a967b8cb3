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
from p2pfl_b200.utils.lockcheck import make_condition, make_lock, make_rlock

_HOST = socket.gethostname()
_PID = os.getpid()

# arenas of nodes living in THIS process: addr -> DeviceArena
_local_arenas: Dict[str, DeviceArena] = {}
_local_lock = make_lock("nvlink.local_arenas")


class ArenaInfoCommand(Command):
    """``arena_info [geo_key, layout_hash, device, pid, host, ipc_hex, n_samples, fused]``.
    A node without a device arena announces too (geometry ``0:0:0:0``), so that its peers know
    rather than wait."""

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
        # willing and able to take part in fused exchanges (has an arena, Settings.FUSED_EXCHANGE)
        self.fused = bool(int(args[7])) if len(args) > 7 else self.geo.n > 0
        self.mapped: Optional[PeerArena] = None

    @property
    def has_arena(self) -> bool:
        return self.geo.n > 0


class NvlinkCommunicationProtocol(BaseProtocol):
    """``Node(model, data, protocol=NvlinkCommunicationProtocol, learner=B200Learner)``."""

    device_payloads = True
    control = "grpc"

    def __init__(self, addr: str = "127.0.0.1") -> None:
        transport: Transport
        if self.control == "grpc":
            transport = GrpcTransport(addr)
        elif self.control == "socket":
            from p2pfl_b200.communication.sockets import SocketTransport

            transport = SocketTransport(addr)
        else:
            transport = MemoryTransport(addr)
        super().__init__(transport)
        self._node: Any = None
        self._peers: Dict[str, PeerInfo] = {}
        self._peers_lock = make_lock("nvlink.peers")
        self._nonce = 0
        self._no_arena: set = set()
        self._my_peerinfo: Optional[PeerInfo] = None
        self._cancelled_round = -1  # highest round whose fused exchange somebody called off
        self._watch_flags: List[Any] = []  # (round, cancel word) of the native waits in progress
        self._announced_to: set = set()  # peers that certainly hold this node's current arena_info
        self.add_command(ArenaInfoCommand(self))
        self.add_command(FusedCancelCommand(self))
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
        self._cancelled_round = -1  # a new experiment starts
        layout_hash = getattr(getattr(learner, "layout", None), "hash", "")
        n_samples = str(learner.get_num_samples()[0])
        if arena is None:
            args = ["0:0:0:0", layout_hash, "-1", str(_PID), _HOST, "", n_samples, "0"]
        else:
            with _local_lock:
                _local_arenas[self.addr] = arena
            learner.plane = self
            try:
                ipc = arena.ipc_handle().hex()
            except Exception as e:  # IPC export can be refused in restricted containers
                logger.debug(self.addr, f"arena IPC export failed: {e}")
                ipc = ""
            args = [arena.geo.key(), layout_hash, str(arena.device), str(_PID), _HOST, ipc, n_samples,
                    str(int(bool(Settings.FUSED_EXCHANGE)))]  # fmt: skip
        self._my_info = args
        self._my_peerinfo = PeerInfo(args)
        # who certainly has this announcement: the extended direct neighbours of this moment (the
        # broadcast is synchronous).  Anybody else is sent a copy right before the first device
        # reference that needs it (``_lower_payload``), on the same ordered link.
        direct = self._neighbors.get_all(only_direct=True)
        self._announced_to = {n for n in direct if n in self._extended}
        self.broadcast(self.build_msg("arena_info", args))

    def _note_extended(self, addr: str) -> None:
        """A peer showed this framework's mark only now: an ``arena_info`` broadcast before that moment
        skipped it (extended commands go to extended peers only), so it gets its copy here."""
        fresh = addr not in self._extended
        super()._note_extended(addr)
        args = getattr(self, "_my_info", None)
        if fresh and args is not None and self._running:
            self._gossiper.add_message(self.build_msg("arena_info", args), [addr])

    def _on_arena_info(self, source: str, args: List[str]) -> None:
        info = PeerInfo(args)
        with self._peers_lock:
            old = self._peers.get(source)
            if old is not None and old.ipc == info.ipc and old.pid == info.pid and old.geo == info.geo:
                old.n_samples, old.layout_hash, old.fused = info.n_samples, info.layout_hash, info.fused
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
            if info is None or info.host != _HOST or not info.has_arena:
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
        if info is None and arena is not None and nei not in self._no_arena and self._node is not None and nei in self._extended:
            # one of ours that has not announced yet is still building its learner (every extended
            # node announces, arena or not): its announcement is milliseconds away
            self._node.state.wait_for(lambda: nei in self._peers, Settings.ARENA_INFO_WAIT)
            with self._peers_lock:
                info = self._peers.get(nei)
            if info is None:
                self._no_arena.add(nei)
        reachable = (
            arena is not None
            and info is not None
            and info.has_arena
            and info.host == _HOST
            and info.layout_hash == learner.layout.hash
            and (info.pid == _PID or bool(getattr(self, "_my_info", ["", "", "", "", "", ""])[5]))
        )
        if not reachable:
            self.stats["host"] += 1
            return Weights(w.source, w.cmd, w.round, learner.encode_parameters(params=payload.params), w.contributors, w.weight)
        mine = getattr(self, "_my_info", None)
        if mine is not None and nei not in self._announced_to:
            # the receiver resolves the reference against MY announcement: make sure it has one (it may
            # have shown up as an extended peer after the broadcast, or be reachable only indirectly)
            self._announced_to.add(nei)
            note = self.build_msg("arena_info", mine)
            note.ttl = 1
            super().send(nei, note, True)
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
    def pull(self, ref: DeviceRef, source: Optional[str] = None) -> Any:
        return self._pull(ref, source)

    def _pull(self, ref: DeviceRef, source: Optional[str] = None) -> Any:
        """Read the sender's tensor over NVLink into a private bf16/fp32 buffer."""
        learner = self._learner()
        if learner is None or getattr(learner, "arena", None) is None:
            raise RuntimeError("no device learner to receive into (learner not built yet)")
        if ref.layout_hash != learner.layout.hash:
            from p2pfl_b200.exceptions import ModelNotMatchingError

            raise ModelNotMatchingError("layout hash differs from the local model")
        from p2pfl_b200.exceptions import DecodingParamsError

        # the reference arrived in an unauthenticated message: nothing in it sizes a read of peer
        # memory before it has been checked against what is known about that peer
        if ref.region not in ("params", "outbox"):
            raise DecodingParamsError(f"device reference names an unknown region {ref.region!r}")
        if source is not None and ref.addr != source:
            raise DecodingParamsError(f"device reference of {ref.addr} arrived in a message from {source}")
        peer = self._map_peer(ref.addr)
        if peer is None:
            # Typically the reference overtook its sender's arena_info (which goes out when the sender
            # learns that this node speaks the extended commands).  No waiting here: with one stream
            # socket per neighbour the announcement queues up BEHIND the very message this thread is
            # handling.  The payload is reported as unavailable, the command drops it, and the
            # sender's gossip loop offers the model again a period later.
            from p2pfl_b200.exceptions import PayloadUnavailableError

            raise PayloadUnavailableError(f"arena of {ref.addr} is not mapped here")
        if int(ref.n) != learner.layout.total or int(ref.n) != peer.geo.n:
            raise DecodingParamsError(f"device reference has {ref.n} elements, the layout has {learner.layout.total}")
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
    def cancel_fused(self, rnd: int) -> None:
        """Tell the round's trainers that this node will NOT take part in (or could not finish) the
        fused exchange of round ``rnd``: whoever has not launched yet skips it, whoever is spinning
        in it aborts, and everybody exchanges through the ordinary add_model path instead."""
        self._note_cancel(rnd)
        try:
            self.broadcast_async(self.build_msg("fused_cancel", [], round=rnd))
        except Exception as e:
            logger.debug(self.addr, f"fused_cancel not sent: {e}")

    def _watch_begin(self, rnd: int) -> Any:
        """A cancel word for one native wait (``ops.rt_watch``) on the exchange of round ``rnd``;
        ``_note_cancel`` sets the words of the rounds it calls off."""
        import ctypes

        flag = ctypes.c_int32(0)
        with self._peers_lock:
            self._watch_flags.append((rnd, flag))
        if self.fused_cancelled(rnd):  # after registering: a cancellation cannot slip between the two
            flag.value = 1
        return flag

    def _watch_end(self, flag: Any) -> None:
        with self._peers_lock:
            self._watch_flags = [(r, f) for r, f in self._watch_flags if f is not flag]

    def _note_cancel(self, rnd: int) -> None:
        if rnd > self._cancelled_round:
            self._cancelled_round = rnd
        with self._peers_lock:
            for r, f in self._watch_flags:  # wake the native watchdogs at once (they poll this word)
                if r <= rnd:
                    f.value = 1
        if self._node is not None:
            st = self._node.state
            if st.fused_complete_round == rnd:
                # a trainer could not finish the exchange this node completed: it will need the
                # aggregate through the ordinary diffusion after all
                st.fused_complete_round = None
            st.notify()

    def fused_cancelled(self, rnd: int) -> bool:
        return self._cancelled_round >= rnd

    def plan_fused_exchange(self, state: Any, learner: Any, wait: Optional[float] = None) -> Optional["FusedExchange"]:
        """A plan for this round's train set, or None (exchange through add_model messages).

        The answer must be the same on every trainer, or one of them would spin on flags nobody
        writes.  It is therefore a function of what every node eventually knows alike - the train
        set and the ``arena_info`` each member announced (geometry, layout, host, process, GPU,
        whether it is willing) - and a member whose announcement has not arrived yet is WAITED for
        (every extended node announces, with or without an arena) rather than guessed.  What can
        still differ between trainers is local: an IPC mapping that fails here.  That trainer then
        says so (``fused_cancel``) and the others fall back with it."""
        rnd = state.round
        arena = getattr(learner, "arena", None)
        train_set = sorted(state.train_set)
        if arena is None or rnd is None or state.addr not in train_set or len(train_set) < 2:
            return None
        others = [n for n in train_set if n != state.addr]
        if self.fused_cancelled(rnd):
            return None
        deadline = time.time() + (Settings.FUSED_INFO_WAIT if wait is None else wait)

        def settled() -> bool:
            if self.fused_cancelled(rnd) or state.round is None:
                return True
            known = self.get_neighbors(only_direct=False)
            with self._peers_lock:
                # a member that left the network, or that is not one of ours (a reference node
                # never announces), settles the question as well: no fused exchange
                return all(n in self._peers or n not in known or (n in known and self._is_plain(n)) for n in others)

        state.wait_for(settled, max(0.0, deadline - time.time()))
        if self.fused_cancelled(rnd) or state.round is None:
            return None
        with self._peers_lock:
            infos = {n: self._peers.get(n) for n in others}
        mine = self._my_peerinfo
        if mine is None or not mine.fused:
            return None
        for n in others:
            info = infos[n]
            if info is None or not info.fused or info.host != _HOST:
                return None
            if info.layout_hash != learner.layout.hash or info.geo != arena.geo:
                return None
            if info.device == arena.device and info.pid != _PID:
                # kernels that wait on one another must not be separate launches on one GPU
                # (B200_PROFILING.md): co-located trainers are grouped into ONE cooperative launch,
                # which only a single process can do
                return None
        if len(train_set) > arena.geo.slots or len(others) > ops.MAX_PEERS:
            return None
        # ---- from here on a failure is local to this node: tell the others
        peers: Dict[str, PeerArena] = {}
        samples: Dict[str, int] = {state.addr: learner.get_num_samples()[0]}
        for n in others:
            p = self._map_peer(n)
            if p is None:
                logger.warning(self.addr, f"fused exchange of round {rnd} cancelled: cannot map the arena of {n}")
                self.cancel_fused(rnd)
                return None
            peers[n] = p
            samples[n] = infos[n].n_samples
        local = [state.addr] + [n for n in others if infos[n].pid == _PID and infos[n].device == arena.device]
        return FusedExchange(self, state, learner, train_set, peers, samples, sorted(local))

    def _is_plain(self, addr: str) -> bool:
        """Known neighbour that does not speak this framework's extra commands."""
        return addr not in self._extended and addr in self._neighbors.get_all(only_direct=True)


class FusedCancelCommand(Command):
    """``fused_cancel`` (round in the message's round field): see ``cancel_fused``."""

    def __init__(self, proto: "NvlinkCommunicationProtocol") -> None:
        self._proto = proto

    @staticmethod
    def get_name() -> str:
        return "fused_cancel"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        self._proto._note_cancel(int(round))


class _DeviceGroup:
    """Rendezvous of the trainers of ONE process that share a GPU in one round: the last to
    arrive launches the single cooperative kernel that runs all their exchanges."""

    def __init__(self, expected: List[str]) -> None:
        self.expected = list(expected)
        self.members: Dict[str, Dict[str, Any]] = {}
        self.cond = make_condition("nvlink.group")
        self.done: Optional[torch.cuda.Event] = None
        self.began: Optional[torch.cuda.Event] = None
        self.failed: Optional[str] = None
        self.keep: List[Any] = []  # device tables must outlive the launch
        self.left = 0


_groups: Dict[Tuple[int, int], _DeviceGroup] = {}
_groups_lock = make_lock("nvlink.groups")
_group_streams: Dict[int, torch.cuda.Stream] = {}


def _group_for(device: int, seq: int, expected: List[str]) -> _DeviceGroup:
    with _groups_lock:
        g = _groups.get((device, seq))
        if g is None:
            g = _groups[(device, seq)] = _DeviceGroup(expected)
        return g


def _group_leave(device: int, seq: int, grp: _DeviceGroup) -> None:
    """A member is done with the group; the last one out removes it from the registry (sequence
    numbers may repeat from one experiment to the next: a finished group must never be found again)."""
    with _groups_lock:
        grp.left += 1
        if grp.left >= len(grp.members) and _groups.get((device, seq)) is grp:
            del _groups[(device, seq)]


def _group_stream(device: torch.device) -> torch.cuda.Stream:
    idx = device.index or 0
    with _groups_lock:
        s = _group_streams.get(idx)
        if s is None:
            s = _group_streams[idx] = torch.cuda.Stream(device=device)
        return s


class FusedExchange:
    """One all-to-all push + FedAvg among the train set: a single kernel per node, or - when
    several trainers of this process share the GPU - a single cooperative kernel for all of them."""

    def __init__(self, proto: NvlinkCommunicationProtocol, state: Any, learner: Any, train_set: List[str],
                 peers: Dict[str, PeerArena], samples: Dict[str, int], local: Optional[List[str]] = None) -> None:  # fmt: skip
        self.proto, self.state, self.learner = proto, state, learner
        self.train_set, self.peers, self.samples = train_set, peers, samples
        self.local = local or [state.addr]
        self.complete = False
        self._launched: Any = None

    def _tables(self, seq: int) -> Tuple[Any, Any]:
        arena = self.learner.arena
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
        return push, red

    @property
    def can_prelaunch(self) -> bool:
        """True when this node's exchange is a launch of its own (no other trainer of this process on
        the GPU to meet first): it can be queued behind the training kernels before they have run."""
        return len(self.local) == 1

    def launch(self, seq: int) -> None:
        """Queue the single-member exchange on the learner's stream; ``finish`` waits for it."""
        learner, arena = self.learner, self.learner.arena
        push, red = self._tables(seq)
        timeout_cycles = int(Settings.AGGREGATION_TIMEOUT * ops.rt_sm_clock_hz(arena.device))
        snap = learner.exchange_snapshot()
        began = torch.cuda.Event(enable_timing=True)
        done = torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(learner.stream):
            arena.status.zero_()
            arena.abort.zero_()
            began.record(learner.stream)
            ops.gossip_fused(arena.params, learner.w16, arena.geo.tile, seq, push, red, arena.abort, timeout_cycles,
                             arena.status, snapshot=snap)  # fmt: skip
            done.record(learner.stream)
        self._launched = (seq, began, done, snap)

    def finish(self) -> Optional[Tuple[List[str], int]]:
        """Wait for a ``launch``-ed exchange; same result as ``run``."""
        seq, began, done, snap = self._launched
        me = self.state.addr
        others = [n for n in self.train_set if n != me]
        total = float(sum(self.samples[n] for n in self.train_set))
        return self._finish(seq, began, done, snap, others, total)

    def run(self, seq: int) -> Optional[Tuple[List[str], int]]:
        """Launch, wait, and return (contributors, total weight) - or None when the exchange did
        not complete: the pre-exchange model is back in place then and the caller exchanges through
        add_model messages (a partly averaged model is never labelled as the round's aggregate)."""
        learner, arena = self.learner, self.learner.arena
        me = self.state.addr
        others = [n for n in self.train_set if n != me]
        total = float(sum(self.samples[n] for n in self.train_set))
        if len(self.local) == 1:
            self.launch(seq)
            return self.finish()
        # several trainers of this process share the GPU: meet them, ONE cooperative launch for all
        push, red = self._tables(seq)
        timeout_cycles = int(Settings.AGGREGATION_TIMEOUT * ops.rt_sm_clock_hz(arena.device))
        snap = learner.exchange_snapshot()
        grp = self._join_group(seq, push, red, snap, timeout_cycles)
        if grp is None:
            return None
        learner.stream.wait_event(grp.done)
        try:
            return self._finish(seq, grp.began, grp.done, snap, others, total)
        finally:
            _group_leave(arena.device, seq, grp)

    def _finish(self, seq: int, began: Any, done: Any, snap: Any, others: List[str], total: float) -> Optional[Tuple[List[str], int]]:
        learner, arena, proto = self.learner, self.learner.arena, self.proto
        me, rnd = self.state.addr, self.state.round
        wire_bf16 = arena.geo.wire_bf16
        # Watchdog (native, csrc/watchdog.cu): this thread waits for the kernel without the interpreter
        # lock; a cancellation (fused_cancel from a peer, any thread) or AGGREGATION_TIMEOUT makes the
        # watchdog flip the abort word itself.  Every WATCH_SLICE seconds it hands back for the checks
        # that need the interpreter: a contributor that left the network, an experiment that stopped.
        cancel = proto._watch_begin(rnd)
        try:
            t0 = time.time()
            while True:
                left = Settings.AGGREGATION_TIMEOUT - (time.time() - t0)
                verdict, _ = ops.rt_watch(arena.device, done, arena.abort, cancel, max(0.0, left), Settings.WATCH_SLICE)
                if verdict != ops.WATCH_SLICE:
                    break
                gone = [n for n in others if n not in proto.get_neighbors(only_direct=False)]
                if gone or self.state.round is None or proto.fused_cancelled(rnd):
                    cancel.value = 1
        finally:
            proto._watch_end(cancel)
        status = arena.status.cpu()
        self.complete = int(status[0]) == 0
        proto.stats["fused_ms"] += began.elapsed_time(done)  # includes the wait for the slowest peer's tiles
        if not self.complete:
            missing = [n for j, n in enumerate(others) if int(status[2]) >> j & 1]
            logger.warning(me, f"fused exchange of round {rnd} incomplete (missing {missing}); restoring the local model")
            with torch.cuda.stream(learner.stream):
                if snap is not None:
                    arena.params.copy_(snap)
                learner._params_written()
            learner.stream.synchronize()
            proto.stats["fused_incomplete"] = proto.stats.get("fused_incomplete", 0) + 1
            proto.cancel_fused(rnd)
            return None
        learner._params_written_by_plane()
        proto.stats["fused"] += 1
        proto.stats["fused_bytes"] += arena.n * (2 if wire_bf16 else 4) * len(others)
        return list(self.train_set), int(total)

    def _join_group(self, seq: int, push: Any, red: Any, snap: Any, timeout_cycles: int) -> Optional[_DeviceGroup]:
        learner, arena, proto = self.learner, self.learner.arena, self.proto
        me, rnd = self.state.addr, self.state.round
        ready = torch.cuda.Event()
        ready.record(learner.stream)
        grp = _group_for(arena.device, seq, self.local)
        member = ops.make_group_member(arena.params, learner.w16, snap, arena.abort, arena.status)
        with grp.cond:
            grp.members[me] = {"push": push, "red": red, "member": member, "ready": ready, "arena": arena}
            if grp.failed is None and len(grp.members) == len(grp.expected) and grp.done is None:
                try:
                    self._launch_group(grp, seq, timeout_cycles)
                except Exception as e:  # a launch failure must not leave the others waiting
                    grp.failed = f"{type(e).__name__}: {e}"
                grp.cond.notify_all()
            else:
                # the other local trainers are still training; a cancellation ends the wait
                deadline = time.time() + Settings.AGGREGATION_TIMEOUT
                while grp.done is None and grp.failed is None:
                    if proto.fused_cancelled(rnd) or self.state.round is None or time.time() > deadline:
                        grp.failed = grp.failed or "cancelled before every local trainer had joined"
                        grp.cond.notify_all()
                        break
                    grp.cond.wait(0.005)
        if grp.failed is not None:
            logger.warning(me, f"group exchange of round {rnd} not launched: {grp.failed}")
            proto.cancel_fused(rnd)
            _group_leave(arena.device, seq, grp)
            return None
        return grp

    def _launch_group(self, grp: _DeviceGroup, seq: int, timeout_cycles: int) -> None:
        """Called by the last local trainer to arrive, with the group's lock held."""
        arena = self.learner.arena
        order = [a for a in grp.expected if a in grp.members]
        gs = _group_stream(self.learner.device)
        with torch.cuda.stream(gs):
            for a in order:
                gs.wait_event(grp.members[a]["ready"])
                grp.members[a]["arena"].status.zero_()
                grp.members[a]["arena"].abort.zero_()
            dev = self.learner.device
            members_dev = ops.tables_to_device([grp.members[a]["member"] for a in order], dev)
            push_dev = ops.tables_to_device([grp.members[a]["push"] for a in order], dev)
            red_dev = ops.tables_to_device([grp.members[a]["red"] for a in order], dev)
            grp.keep = [members_dev, push_dev, red_dev]
            began = torch.cuda.Event(enable_timing=True)
            began.record(gs)
            ops.gossip_fused_group(arena.device, arena.n, arena.geo.tile, seq, members_dev, push_dev, red_dev, len(order), timeout_cycles)
            done = torch.cuda.Event(enable_timing=True)
            done.record(gs)
        grp.began, grp.done = began, done


class NvlinkSocketProtocol(NvlinkCommunicationProtocol):
    """Same data plane; control plane over framed stream sockets (one process per GPU, all nodes
    of this framework: several times cheaper per message than gRPC)."""

    control = "socket"


class NvlinkMemoryProtocol(NvlinkCommunicationProtocol):
    """Same data plane, in-process control plane (several nodes per process / per GPU)."""

    control = "memory"
